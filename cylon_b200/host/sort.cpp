// sort.cpp -- cylon::Sort, MapToSortPartitions and DistributedSort over the kernels of csrc/sort.cu.
//
// Reference: table.cpp:291-382 (Sort, DistributedSort), partition/partition.cpp:168-198 (MapToSortPartitions),
// arrow/arrow_partition_kernels.cpp:334-455 (RangePartitionKernel). The sequence is the reference's: range-partition on the first
// sort column -> Split -> all-to-all -> local sort by all sort columns. Differences that keep the result a valid distributed
// sort: the sample is a fixed stride instead of random; total_samples is the number of values actually sampled over all ranks (the
// reference multiplies its OWN num_samples by the world size, the same number whenever the ranks hold equal row counts); null bits
// survive the shuffle (the reference's Split drops them).
#include <cuda_runtime.h>

#include <algorithm>

#include "cylon/partition/partition.hpp"
#include "cylon/table.hpp"
#include "table_ops.hpp"

namespace cylon {

namespace {

using detail::DevBuf;

Status CheckSortColumn(const Table &t, int32_t col, const char *who) {
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(t, {col}, who));
  const auto &c = t.GetColumn(col);
  if (c->is_var_width() || c->byte_width() > 8 || c->type() == CYL_FIXED_SIZE_BINARY || c->type() == CYL_BOOL)
    return Status(Code::NotImplemented, std::string(who) + ": sort columns must be numeric or temporal");
  return Status::OK();
}

// device form of MapToSortPartitions: pid (uint32 per row) and, when asked for, the rows per partition, both in `stream` order
Status RangePartitionIds(const std::shared_ptr<Table> &table, int32_t column_idx, uint32_t num_partitions, bool ascending, uint64_t num_samples,
                         uint32_t num_bins, uint32_t *pid, uint64_t *partition_hist) {
  auto ctx = table->GetContext();
  void *stream = ctx->GetStream();
  const int64_t n = table->Rows();
  const cyl_column_t col = table->GetColumn(column_idx)->desc();
  if (num_samples == 0) num_samples = (uint64_t)std::min((int64_t)((double)n * 0.01), n);  // partition.cpp:181
  if (num_bins == 0) num_bins = num_partitions * 16;                                         // partition.cpp:182
  const int64_t samples = (int64_t)std::min<uint64_t>(num_samples, (uint64_t)n);
  const int world = ctx->GetWorldSize();
  auto comm = ctx->GetCommunicator();

  // min / max of the sample over all ranks (compute::MinMax, arrow_partition_kernels.cpp:384-390), as order keys
  uint64_t min_key = ~0ull, max_key = 0;
  int64_t sampled = 0;
  CYL_RETURN_IF_FAILED(cyl_range_sample_minmax(&col, n, samples, &min_key, &max_key, &sampled, stream));
  int64_t total_sampled = sampled;
  if (world > 1) {
    const int64_t mine[3] = {(int64_t)min_key, (int64_t)max_key, sampled};
    std::vector<int64_t> all((size_t)3 * world);
    RETURN_CYLON_STATUS_IF_FAILED(comm->AllGather(mine, 3, all.data(), stream));
    total_sampled = 0;
    for (int r = 0; r < world; r++) {
      if (all[(size_t)3 * r + 2] == 0) continue;  // that rank sampled nothing
      min_key = std::min(min_key, (uint64_t)all[(size_t)3 * r]);
      max_key = std::max(max_key, (uint64_t)all[(size_t)3 * r + 1]);
      total_sampled += all[(size_t)3 * r + 2];
    }
  }

  // histogram of the samples over num_bins + 2 bins, summed over the ranks (mpi::AllReduce, :392-407)
  const size_t bins = (size_t)num_bins + 2;
  std::vector<uint64_t> counts(bins, 0);
  if (total_sampled > 0) {
    CYL_RETURN_IF_FAILED(cyl_range_sample_histogram(&col, n, samples, min_key, max_key, num_bins, counts.data(), stream));
    if (world > 1) {
      std::vector<int64_t> mine(counts.begin(), counts.end()), all(bins * (size_t)world);
      RETURN_CYLON_STATUS_IF_FAILED(comm->AllGather(mine.data(), (int)bins, all.data(), stream));
      std::fill(counts.begin(), counts.end(), 0);
      for (int r = 0; r < world; r++)
        for (size_t b = 0; b < bins; b++) counts[b] += (uint64_t)all[(size_t)r * bins + b];
    }
  }
  std::vector<uint32_t> bin_to_partition(bins, 0);  // quantile split of the bins, in float like the reference (:409-427)
  CYL_RETURN_IF_FAILED(cyl_range_quantile_split(counts.data(), num_bins, (uint64_t)total_sampled, num_partitions, bin_to_partition.data()));
  CYL_RETURN_IF_FAILED(cyl_range_partition(&col, n, min_key, max_key, total_sampled, num_bins, bin_to_partition.data(), num_partitions, ascending ? 1 : 0,
                                           pid, partition_hist, stream));
  return Status::OK();
}

}  // namespace

// ---------------------------------------------------------------------------------------------- local sort
Status Sort(std::shared_ptr<Table> &table, const std::vector<int32_t> &sort_columns, std::shared_ptr<Table> &out,
            const std::vector<bool> &sort_direction) {  // table.cpp:311-328
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  if (sort_columns.empty()) return Status(Code::Invalid, "Sort: no sort column");
  if (sort_columns.size() != sort_direction.size())
    return Status(Code::Invalid, "No of sort columns and no of sort direction indicators mismatch");  // arrow_kernels.cpp:417-420
  for (int32_t c : sort_columns) RETURN_CYLON_STATUS_IF_FAILED(CheckSortColumn(*table, c, "Sort"));
  auto ctx = table->GetContext();
  void *stream = ctx->GetStream();
  const int64_t n = table->Rows();
  const auto keys = detail::Descs(*table, std::vector<int>(sort_columns.begin(), sort_columns.end()));
  std::vector<int32_t> asc;
  for (bool d : sort_direction) asc.push_back(d ? 1 : 0);
  DevBuf idx;
  RETURN_CYLON_STATUS_IF_FAILED(idx.alloc((size_t)(n + 1) * 8, stream));
  CYL_RETURN_IF_FAILED(cyl_sort_indices(keys.data(), (int32_t)keys.size(), asc.data(), n, idx.as<int64_t>(), stream));
  std::vector<std::shared_ptr<Column>> cols;
  // arrow Take (util/arrow_utils.cpp:56-64) carries the source validity, unlike the join's copy_array_by_indices
  RETURN_CYLON_STATUS_IF_FAILED(detail::GatherColumns(*table, idx.as<int64_t>(), n, false, true, "", stream, &cols));
  out = std::make_shared<Table>(std::move(cols), ctx);
  return Status::OK();
}

Status Sort(std::shared_ptr<Table> &table, int sort_column, std::shared_ptr<Table> &output, bool ascending) {  // table.cpp:291-303
  return Sort(table, std::vector<int32_t>{sort_column}, output, std::vector<bool>{ascending});
}

Status Sort(std::shared_ptr<Table> &table, const std::vector<int32_t> &sort_columns, std::shared_ptr<Table> &out, bool ascending) {  // table.cpp:305-309
  return Sort(table, sort_columns, out, std::vector<bool>(sort_columns.size(), ascending));
}

// ---------------------------------------------------------------------------------------------- range partitioning
Status MapToSortPartitions(const std::shared_ptr<Table> &table, int32_t column_idx, uint32_t num_partitions, uint32_t **target_partitions,
                           uint64_t **partition_hist, bool ascending, uint64_t num_samples, uint32_t num_bins) {
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  if (num_partitions < 1) return Status(Code::Invalid, "num_partitions must be positive");
  RETURN_CYLON_STATUS_IF_FAILED(CheckSortColumn(*table, column_idx, "MapToSortPartitions"));
  void *pid = nullptr, *hist = nullptr;
  CYL_RETURN_IF_FAILED(cyl_malloc(&pid, (table->Rows() + 1) * 4));
  int rc = cyl_malloc(&hist, (int64_t)num_partitions * 8);
  if (rc != CYL_OK) {
    cyl_free(pid);
    return detail::FromCode(rc);
  }
  Status st = RangePartitionIds(table, column_idx, num_partitions, ascending, num_samples, num_bins, static_cast<uint32_t *>(pid),
                                static_cast<uint64_t *>(hist));
  if (!st.is_ok()) {
    cyl_free(pid);
    cyl_free(hist);
    return st;
  }
  *target_partitions = static_cast<uint32_t *>(pid);
  *partition_hist = static_cast<uint64_t *>(hist);
  return Status::OK();
}

Status MapToSortPartitions(const std::shared_ptr<Table> &table, int32_t column_idx, uint32_t num_partitions, std::vector<uint32_t> &target_partitions,
                           std::vector<uint32_t> &partition_hist, bool ascending, uint64_t num_samples, uint32_t num_bins) {
  uint32_t *pid = nullptr;
  uint64_t *hist = nullptr;
  RETURN_CYLON_STATUS_IF_FAILED(MapToSortPartitions(table, column_idx, num_partitions, &pid, &hist, ascending, num_samples, num_bins));
  cudaStream_t s = static_cast<cudaStream_t>(table->GetContext()->GetStream());
  const int64_t n = table->Rows();
  target_partitions.assign((size_t)n, 0);
  std::vector<uint64_t> h64(num_partitions, 0);
  cudaError_t e = cudaSuccess;
  if (n > 0) e = cudaMemcpyAsync(target_partitions.data(), pid, (size_t)n * 4, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(h64.data(), hist, (size_t)num_partitions * 8, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  cyl_free(pid);
  cyl_free(hist);
  if (e != cudaSuccess) return Status(Code::ExecutionError, std::string("MapToSortPartitions: ") + cudaGetErrorString(e));
  partition_hist.assign(num_partitions, 0);
  for (uint32_t p = 0; p < num_partitions; p++) partition_hist[p] = (uint32_t)h64[p];  // the reference's counters are 32-bit
  return Status::OK();
}

// ---------------------------------------------------------------------------------------------- distributed sort
Status DistributedSort(std::shared_ptr<Table> &table, const std::vector<int> &sort_columns, std::shared_ptr<Table> &output,
                       const std::vector<bool> &sort_direction, SortOptions sort_options) {  // table.cpp:338-382
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  if (sort_columns.empty()) return Status(Code::Invalid, "DistributedSort: no sort column");
  if (sort_columns.size() != sort_direction.size())
    return Status(Code::Invalid, "No of sort columns and no of sort direction indicators mismatch");
  for (int c : sort_columns) RETURN_CYLON_STATUS_IF_FAILED(CheckSortColumn(*table, c, "DistributedSort"));
  auto ctx = table->GetContext();
  const int world = ctx->GetWorldSize();
  const std::vector<int32_t> cols32(sort_columns.begin(), sort_columns.end());
  std::shared_ptr<Table> local = table;
  if (world > 1) {
    detail::StageTrace trace("DistributedSort", ctx->GetStream());
    DevBuf pid;
    RETURN_CYLON_STATUS_IF_FAILED(pid.alloc((size_t)(table->Rows() + 1) * 4, ctx->GetStream()));
    RETURN_CYLON_STATUS_IF_FAILED(RangePartitionIds(table, sort_columns[0], (uint32_t)world, sort_direction[0], sort_options.num_samples,
                                                    sort_options.num_bins, pid.as<uint32_t>(), nullptr));
    trace.mark("range partition");
    std::shared_ptr<Table> shuffled;
    RETURN_CYLON_STATUS_IF_FAILED(detail::ShuffleByPartitionIds(ctx, table, pid.as<uint32_t>(), shuffled));
    trace.mark("shuffle");
    local = shuffled;
  }
  return Sort(local, cols32, output, sort_direction);
}

Status DistributedSort(std::shared_ptr<Table> &table, int sort_column, std::shared_ptr<Table> &output, bool ascending, SortOptions sort_options) {
  return DistributedSort(table, std::vector<int>{sort_column}, output, std::vector<bool>{ascending}, sort_options);  // table.cpp:330-336
}

}  // namespace cylon
