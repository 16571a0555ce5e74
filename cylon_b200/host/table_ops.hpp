// table_ops.hpp -- internal building blocks shared by the host operators (not installed).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "cylon/table.hpp"
#include "util.hpp"

namespace cylon {
namespace detail {

std::vector<cyl_column_t> Descs(const Table &t, const std::vector<int> &cols);
std::vector<cyl_column_t> AllDescs(const Table &t);
Status CheckColumns(const Table &t, const std::vector<int> &cols, const char *what);

// copy_array_by_indices for every column of `t` (util/copy_arrray.cpp:142-292): out[i] = idx[i] < 0 ? null : t[idx[i]].
// may_be_negative: idx can hold -1 (then a validity bitmap is produced); source nulls are always kept.
Status GatherColumns(const Table &t, const int64_t *idx, int64_t m, bool may_be_negative, bool keep_source_validity, const std::string &prefix,
                     void *stream,
                     std::vector<std::shared_ptr<Column>> *out);
// arrow::compute::Filter with a byte mask
Status FilterTable(const std::shared_ptr<Table> &t, const uint8_t *mask, std::shared_ptr<Table> *out);
// arrow::ConcatenateTables + CombineChunks of two tables with the same schema
Status ConcatTables(const std::shared_ptr<Table> &a, const std::shared_ptr<Table> &b, std::shared_ptr<Table> *out);
Status VerifyTableSchema(const Table &a, const Table &b);
// all-to-all of a table whose rows already know their target rank (device uint32 per row): Split + exchange, chunks ordered by
// source rank, row order inside a chunk = input order (host/shuffle.cpp)
Status ShuffleByPartitionIds(std::shared_ptr<CylonContext> &ctx, std::shared_ptr<Table> &table, const uint32_t *target_ranks, std::shared_ptr<Table> &out);

}  // namespace detail
}  // namespace cylon
