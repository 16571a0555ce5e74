// DistributedHashGroupBy / HashGroupBy with the reference's signatures
// (cpp/src/cylon/groupby/groupby.hpp:23-33, groupby/hash_groupby.hpp:31-66).
#pragma once
#include <memory>
#include <utility>
#include <vector>

#include "../compute/aggregate_kernels.hpp"
#include "../table.hpp"

namespace cylon {

Status HashGroupBy(const std::shared_ptr<Table> &table, const std::vector<int32_t> &idx_cols,
                   const std::vector<std::pair<int32_t, compute::AggregationOpId>> &aggregations, std::shared_ptr<Table> &output);
Status HashGroupBy(const std::shared_ptr<Table> &table, int32_t idx_col, const std::vector<int32_t> &aggregate_cols,
                   const std::vector<compute::AggregationOpId> &aggregate_ops, std::shared_ptr<Table> &output);
Status DistributedHashGroupBy(std::shared_ptr<Table> &table, const std::vector<int32_t> &index_cols, const std::vector<int32_t> &aggregate_cols,
                              const std::vector<compute::AggregationOpId> &aggregate_ops, std::shared_ptr<Table> &output);
Status DistributedHashGroupBy(std::shared_ptr<Table> &table, int32_t index_col, const std::vector<int32_t> &aggregate_cols,
                              const std::vector<compute::AggregationOpId> &aggregate_ops, std::shared_ptr<Table> &output);

}  // namespace cylon
