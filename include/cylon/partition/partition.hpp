// MapToHashPartitions / Split with the reference's names (cpp/src/cylon/partition/partition.hpp:32-42,76-85).
// The partition ids and the histogram stay in HBM (device pointers) instead of std::vector<uint32_t>; the histogram
// is 64-bit.
#pragma once
#include <memory>
#include <vector>

#include "../table.hpp"

namespace cylon {

// target_partitions: device uint32[rows]; partition_hist: device uint64[num_partitions]. Both allocated by the callee
// (cyl_malloc'ed; release with cyl_free).
Status MapToHashPartitions(const std::shared_ptr<Table> &table, const std::vector<int32_t> &hash_column_idx, uint32_t num_partitions,
                           uint32_t **target_partitions, uint64_t **partition_hist);
Status MapToHashPartitions(const std::shared_ptr<Table> &table, int32_t hash_column_idx, uint32_t num_partitions,
                           uint32_t **target_partitions, uint64_t **partition_hist);
// stable split by device partition ids into num_partitions tables (row order inside a partition = input order)
Status Split(const std::shared_ptr<Table> &table, uint32_t num_partitions, const uint32_t *target_partitions,
             std::vector<std::shared_ptr<Table>> &output);

}  // namespace cylon
