// MapToHashPartitions / Split / PartitionByHashing with the reference's names (cpp/src/cylon/partition/partition.hpp:32-90).
//
// Two forms of each. The std::vector forms have the reference's signatures (ids and histogram in host vectors) so that
// code written against the reference compiles unchanged; they wrap the device forms and copy the ids / histogram over
// PCIe, so they are for inspection and tests. The device forms keep the partition ids and the (64-bit) histogram in HBM
// and are what the operators use. Split returns cylon::Table partitions (device-resident) where the reference returns
// arrow::Table (host): move them out with Table::ToArrowTable (include/cylon/arrow_bridge.hpp) when host tables are wanted.
#pragma once
#include <memory>
#include <vector>

#include "../table.hpp"

namespace cylon {

// ---- reference signatures (partition/partition.hpp:32-42,76-90)
Status MapToHashPartitions(const std::shared_ptr<Table> &table, int32_t hash_column_idx, uint32_t num_partitions,
                           std::vector<uint32_t> &target_partitions, std::vector<uint32_t> &partition_histogram);
Status MapToHashPartitions(const std::shared_ptr<Table> &table, const std::vector<int32_t> &hash_column_idx, uint32_t num_partitions,
                           std::vector<uint32_t> &target_partitions, std::vector<uint32_t> &partition_histogram);
Status Split(const std::shared_ptr<Table> &table, uint32_t num_partitions, const std::vector<uint32_t> &target_partitions,
             std::vector<std::shared_ptr<Table>> &output);
// partition_hist_ptr only saves the reference a counting pass; here the counts come out of the scatter map anyway. It is
// checked against them (Code::Invalid on a mismatch, where the reference would write past its builders).
Status Split(const std::shared_ptr<Table> &table, uint32_t num_partitions, const std::vector<uint32_t> &target_partitions,
             const std::vector<uint32_t> &partition_hist_ptr, std::vector<std::shared_ptr<Table>> &output);
Status PartitionByHashing(const std::shared_ptr<Table> &table, const std::vector<int32_t> &hash_cols, uint32_t num_partitions,
                          std::vector<std::shared_ptr<Table>> &partitions);

// Range partitioning for DistributedSort (partition/partition.hpp:60-74; partition.cpp:168-198): target partition of every row
// by the value of one numeric column. COLLECTIVE when the table's context is distributed (min/max and the sample histogram are
// reduced over the ranks). num_samples == 0: 1 % of the rows; num_bins == 0: 16 per partition. The sample is rows
// floor(i * rows / num_samples) (the reference draws it at random unless num_samples == rows).
Status MapToSortPartitions(const std::shared_ptr<Table> &table, int32_t column_idx, uint32_t num_partitions, std::vector<uint32_t> &target_partitions,
                           std::vector<uint32_t> &partition_hist, bool ascending = true, uint64_t num_samples = 0, uint32_t num_bins = 0);

// ---- device forms. target_partitions: device uint32[rows]; partition_hist: device uint64[num_partitions]. Both are
// allocated by the callee (cyl_malloc'ed; release with cyl_free).
Status MapToHashPartitions(const std::shared_ptr<Table> &table, const std::vector<int32_t> &hash_column_idx, uint32_t num_partitions,
                           uint32_t **target_partitions, uint64_t **partition_hist);
Status MapToHashPartitions(const std::shared_ptr<Table> &table, int32_t hash_column_idx, uint32_t num_partitions,
                           uint32_t **target_partitions, uint64_t **partition_hist);
Status MapToSortPartitions(const std::shared_ptr<Table> &table, int32_t column_idx, uint32_t num_partitions, uint32_t **target_partitions,
                           uint64_t **partition_hist, bool ascending = true, uint64_t num_samples = 0, uint32_t num_bins = 0);
// stable split by device partition ids into num_partitions tables (row order inside a partition = input order)
Status Split(const std::shared_ptr<Table> &table, uint32_t num_partitions, const uint32_t *target_partitions,
             std::vector<std::shared_ptr<Table>> &output);

}  // namespace cylon
