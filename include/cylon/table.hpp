// cylon::Table and the relational free functions of the hot path, with the reference's names, argument order and
// Status convention (cpp/src/cylon/table.hpp:76-174,258-348,433-436). A Table here is a set of device-resident
// columns (include/cylon/column.hpp) instead of an arrow::Table in host memory; every operator runs as CUDA
// kernels on the context's stream through the C-ABI of include/cylon_b200.h, and the Distributed* variants shuffle
// over NCCL. There is no CPU path.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "column.hpp"
#include "ctx/cylon_context.hpp"
#include "join/join_config.hpp"
#include "status.hpp"

namespace arrow {
class Table;
}

namespace cylon {

class Table {
 public:
  Table(std::vector<std::shared_ptr<Column>> columns, std::shared_ptr<CylonContext> ctx) : columns_(std::move(columns)), ctx_(std::move(ctx)) {}

  // reference: Table::FromColumns (table.hpp:100) / FromArrowTable (table.hpp:84) -- here from device columns ...
  static Status FromColumns(std::shared_ptr<CylonContext> &ctx, std::vector<std::shared_ptr<Column>> columns, std::shared_ptr<Table> &out);
  // ... or from host Arrow-layout buffers (copied to HBM on the context's stream)
  static Status FromHostColumns(std::shared_ptr<CylonContext> &ctx, const std::vector<cyl_column_t> &host_columns,
                                const std::vector<std::string> &names, std::shared_ptr<Table> &out);

  // Arrow host tables in and out (table.hpp:76,96,146 of the reference). Declared here, defined inline in
  // <cylon/arrow_bridge.hpp>, which needs the Arrow C++ headers: include it where these are called.
  static Status FromArrowTable(std::shared_ptr<CylonContext> &ctx, std::shared_ptr<arrow::Table> &table, std::shared_ptr<Table> &out);
  Status ToArrowTable(std::shared_ptr<arrow::Table> &output);
  std::shared_ptr<arrow::Table> get_table();

  int64_t Rows() const { return columns_.empty() ? 0 : columns_[0]->length(); }  // table.hpp:146
  int32_t Columns() const { return (int32_t)columns_.size(); }                   // table.hpp:140
  std::vector<std::string> ColumnNames() const;
  const std::shared_ptr<Column> &GetColumn(int i) const { return columns_[i]; }
  const std::vector<std::shared_ptr<Column>> &GetColumns() const { return columns_; }
  std::shared_ptr<CylonContext> GetContext() const { return ctx_; }               // table.hpp:162
  void retainMemory(bool retain) { retain_ = retain; }                            // table.hpp:168
  bool IsRetain() const { return retain_; }                                       // table.hpp:173
  Status Project(const std::vector<int32_t> &cols, std::shared_ptr<Table> &out) const;  // table.hpp:190 (zero copy)

 private:
  std::vector<std::shared_ptr<Column>> columns_;
  std::shared_ptr<CylonContext> ctx_;
  bool retain_ = true;
};

// --- joins (table.hpp:258,269; table.cpp:428-502) -------------------------------------------------------------
Status Join(std::shared_ptr<Table> &left, std::shared_ptr<Table> &right, const join::config::JoinConfig &join_config,
            std::shared_ptr<Table> &output);
Status DistributedJoin(std::shared_ptr<Table> &left, std::shared_ptr<Table> &right, const join::config::JoinConfig &join_config,
                       std::shared_ptr<Table> &output);

// --- set operations (table.hpp:279-329; table.cpp:531-785) -------------------------------------------------------
Status Union(std::shared_ptr<Table> &first, std::shared_ptr<Table> &second, std::shared_ptr<Table> &output);
Status Subtract(std::shared_ptr<Table> &first, std::shared_ptr<Table> &second, std::shared_ptr<Table> &output);
Status Intersect(std::shared_ptr<Table> &first, std::shared_ptr<Table> &second, std::shared_ptr<Table> &output);
Status DistributedUnion(std::shared_ptr<Table> &left, std::shared_ptr<Table> &right, std::shared_ptr<Table> &output);
Status DistributedSubtract(std::shared_ptr<Table> &left, std::shared_ptr<Table> &right, std::shared_ptr<Table> &output);
Status DistributedIntersect(std::shared_ptr<Table> &left, std::shared_ptr<Table> &right, std::shared_ptr<Table> &output);

// --- shuffle / partition (table.hpp:339,348; table.cpp:133-157,910-921) ------------------------------------------
Status Shuffle(std::shared_ptr<Table> &table, const std::vector<int> &hash_columns, std::shared_ptr<Table> &output);
Status HashPartition(std::shared_ptr<Table> &table, const std::vector<int> &hash_columns, int no_of_partitions,
                     std::unordered_map<int, std::shared_ptr<Table>> *output);

// --- unique (table.hpp:433,436; table.cpp:923-999). Only first == true is supported. -------------
Status Unique(std::shared_ptr<Table> &in, const std::vector<int> &cols, std::shared_ptr<Table> &out, bool first = true);
Status DistributedUnique(std::shared_ptr<Table> &in, const std::vector<int> &cols, std::shared_ptr<Table> &out);

// --- sort (table.hpp:353-404; table.cpp:291-382). Sort columns: numeric / temporal. Ties keep their input order; floats order
// by IEEE total order; validity is carried along and not looked at (as the reference's sorters). ----------------------------
Status Sort(std::shared_ptr<Table> &table, int sort_column, std::shared_ptr<Table> &output, bool ascending = true);
Status Sort(std::shared_ptr<Table> &table, const std::vector<int32_t> &sort_columns, std::shared_ptr<Table> &out, bool ascending);
Status Sort(std::shared_ptr<Table> &table, const std::vector<int32_t> &sort_columns, std::shared_ptr<Table> &out, const std::vector<bool> &sort_direction);
struct SortOptions {  // 0 = default: num_bins = 16 per rank, num_samples = 1 % of the rank's rows (partition.cpp:181-182)
  uint32_t num_bins;
  uint64_t num_samples;
  static SortOptions Defaults() { return {0, 0}; }
};
// Range-partitions on the FIRST sort column (sample -> min/max and bin histogram over all ranks -> quantile split), shuffles, sorts
// locally by all sort columns: rank r holds a sorted run and every value of rank r orders before (ascending) those of rank r + 1.
Status DistributedSort(std::shared_ptr<Table> &table, int sort_column, std::shared_ptr<Table> &output, bool ascending = true,
                       SortOptions sort_options = SortOptions::Defaults());
Status DistributedSort(std::shared_ptr<Table> &table, const std::vector<int> &sort_columns, std::shared_ptr<Table> &output,
                       const std::vector<bool> &sort_direction, SortOptions sort_options = SortOptions::Defaults());

// order-independent checksum of a table (no counterpart in the reference; used to compare full-size results)
Status TableChecksum(const std::shared_ptr<Table> &table, uint64_t *checksum);

}  // namespace cylon
