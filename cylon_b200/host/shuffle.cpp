// shuffle.cpp -- hash partition -> all-to-all -> contiguous table, and the Distributed* operators built on it.
//
// Replaces (reference, relative to cpp/src/cylon/): table.cpp:67-131 (all_to_all_arrow_tables), :133-179
// (shuffle_table_by_hashing / shuffle_two_tables_by_hashing), :482-502 (DistributedJoin), :727-785 (distributed set
// operations), :910-921 (Shuffle), :984-999 (DistributedUnique), partition/partition.cpp:66-217 (Split,
// MapToHashPartitions, PartitionByHashing) and arrow/arrow_all_to_all.cpp:24-268 (per-buffer framing).
//
// What is different by design:
//  * the table is scattered ONCE into a partition-ordered copy (partition p's rows are contiguous, in input order);
//    that copy is the send buffer -- no per-partition arrow::Table objects, no builders;
//  * one AllGather of the P x (rows, bytes per variable-width column, has-nulls per column) count vector replaces the
//    per-buffer header/FIN handshake; every receive offset is known before any payload moves, so each peer's slice
//    lands directly at its final position in the output columns -- the reference's ConcatenateTables + CombineChunks
//    copy (table.cpp:119-128) does not exist;
//  * all column buffers of the table travel in one grouped ncclSend/ncclRecv; sizes are 64-bit (the reference is
//    limited to 2 GiB per buffer per peer by `int` lengths, net/TxRequest.hpp:25);
//  * received chunks are ordered by source rank (the reference orders by arrival: non-deterministic).
#include <cuda_runtime.h>

#include <algorithm>

#include "cylon/partition/partition.hpp"
#include "cylon/table.hpp"
#include "table_ops.hpp"

namespace cylon {

namespace {

using detail::DevBuf;

struct Partitioned {
  std::vector<std::shared_ptr<Column>> cols;  // partition-ordered copy
  std::vector<int64_t> row_off;               // [P+1]
  std::vector<std::vector<int64_t>> byte_off; // per var-width column: [P+1] byte offsets at partition boundaries (else empty)
};

// byte_off[c][p] = offsets of variable-width column c at the first row of partition p (and at the end for p == P)
Status BoundaryByteOffsets(const std::vector<std::shared_ptr<Column>> &cols, const int64_t *part_offsets_dev, uint32_t P, void *stream,
                           std::vector<std::vector<int64_t>> *byte_off) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int nvar = 0;
  for (const auto &c : cols) nvar += c->is_var_width() ? 1 : 0;
  if (nvar == 0) return Status::OK();
  DevBuf picked;
  RETURN_CYLON_STATUS_IF_FAILED(picked.alloc((size_t)nvar * (P + 1) * 4, stream));
  int v = 0;
  for (const auto &c : cols) {
    if (!c->is_var_width()) continue;
    CYL_RETURN_IF_FAILED(cyl_gather_offsets_at(c->desc().offsets, part_offsets_dev, (int64_t)P + 1, picked.as<int32_t>() + (size_t)v * (P + 1), stream));
    v++;
  }
  std::vector<int32_t> host((size_t)nvar * (P + 1), 0);
  CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(host.data(), picked.p, host.size() * 4, cudaMemcpyDeviceToHost, s));
  CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(s));
  v = 0;
  for (size_t c = 0; c < cols.size(); c++) {
    if (!cols[c]->is_var_width()) continue;
    (*byte_off)[c].assign(host.begin() + (size_t)v * (P + 1), host.begin() + (size_t)(v + 1) * (P + 1));
    v++;
  }
  return Status::OK();
}

// MapToHashPartitions + Split fused: returns the partition-ordered table and the partition boundaries (host)
Status PartitionTable(const std::shared_ptr<Table> &table, const std::vector<int> &hash_cols, uint32_t P, Partitioned *out) {
  auto ctx = table->GetContext();
  void *stream = ctx->GetStream();
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t n = table->Rows();
  const int ncols = table->Columns();
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*table, hash_cols, "partition"));

  bool simple = true;  // all fixed width of a vector-friendly size, no validity: one fused kernel pass
  for (const auto &c : table->GetColumns()) {
    const int w = c->byte_width();
    if (c->is_var_width() || c->has_validity() || !(w == 1 || w == 2 || w == 4 || w == 8 || w == 16)) simple = false;
  }
  DevBuf part_offsets;
  RETURN_CYLON_STATUS_IF_FAILED(part_offsets.alloc((P + 1) * sizeof(int64_t), stream));
  out->cols.clear();
  out->byte_off.assign((size_t)ncols, {});
  for (const auto &c : table->GetColumns()) {
    std::shared_ptr<Column> d;
    RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(c->name(), c->type(), c->byte_width(), n, c->data_bytes(), c->has_validity(), stream, &d));
    out->cols.push_back(d);
  }
  const auto all = detail::AllDescs(*table);
  if (simple) {
    std::vector<void *> dst;
    for (auto &c : out->cols) dst.push_back(c->mutable_data());
    std::vector<int32_t> keys(hash_cols.begin(), hash_cols.end());
    CYL_RETURN_IF_FAILED(cyl_partition_scatter_fixed(all.data(), ncols, keys.data(), (int32_t)keys.size(), n, P, dst.data(),
                                                     part_offsets.as<int64_t>(), stream));
  } else {
    DevBuf dest;
    RETURN_CYLON_STATUS_IF_FAILED(dest.alloc((size_t)(n + 1) * 8, stream));
    const auto kd = detail::Descs(*table, hash_cols);
    CYL_RETURN_IF_FAILED(cyl_partition_map(kd.data(), (int32_t)kd.size(), n, P, nullptr, dest.as<int64_t>(), part_offsets.as<int64_t>(), stream));
    for (int c = 0; c < ncols; c++) {
      const cyl_column_t &src = all[c];
      auto &d = out->cols[c];
      if (d->is_var_width()) {
        CYL_RETURN_IF_FAILED(cyl_scatter_varwidth(src.offsets, static_cast<const uint8_t *>(src.data), dest.as<int64_t>(), n,
                                                  d->mutable_offsets(), static_cast<uint8_t *>(d->mutable_data()), stream));
      } else {
        CYL_RETURN_IF_FAILED(cyl_scatter_fixed(src.data, src.byte_width, dest.as<int64_t>(), n, d->mutable_data(), stream));
      }
      if (src.validity) CYL_RETURN_IF_FAILED(cyl_scatter_validity(src.validity, dest.as<int64_t>(), n, d->mutable_validity(), stream));
    }
  }
  out->row_off.assign(P + 1, 0);
  CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(out->row_off.data(), part_offsets.p, (P + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(s));
  // byte offsets of variable-width columns at the partition boundaries: one small gather on the device and one copy per column
  RETURN_CYLON_STATUS_IF_FAILED(BoundaryByteOffsets(out->cols, part_offsets.as<int64_t>(), P, stream, &out->byte_off));
  return Status::OK();
}

// zero-copy slice of rows [r0, r1) of a partition-ordered table (fixed width, byte-aligned validity only when r0 % 8 == 0)
// one byte per row of every column that has a validity bitmap, made once and shared by all partitions' slices
Status UnpackValidity(const Partitioned &pt, void *stream, std::vector<std::unique_ptr<DevBuf>> *bytes) {
  bytes->clear();
  for (const auto &src : pt.cols) {
    bytes->emplace_back(nullptr);
    const cyl_column_t &d = src->desc();
    if (!d.validity) continue;
    bytes->back().reset(new DevBuf());
    RETURN_CYLON_STATUS_IF_FAILED(bytes->back()->alloc((size_t)src->length() + 1, stream));
    CYL_RETURN_IF_FAILED(cyl_unpack_validity(d.validity, src->length(), bytes->back()->as<uint8_t>(), stream));
  }
  return Status::OK();
}

Status SliceRows(const Partitioned &pt, const std::vector<std::unique_ptr<DevBuf>> &validity_bytes, const std::shared_ptr<CylonContext> &ctx, uint32_t p,
                 std::shared_ptr<Table> *out) {
  void *stream = ctx->GetStream();
  const int64_t r0 = pt.row_off[p], r1 = pt.row_off[p + 1], m = r1 - r0;
  std::vector<std::shared_ptr<Column>> cols;
  for (size_t c = 0; c < pt.cols.size(); c++) {
    const auto &src = pt.cols[c];
    const cyl_column_t &d = src->desc();
    std::shared_ptr<Column> col;
    if (src->is_var_width()) {
      const int64_t b0 = pt.byte_off[c][p], b1 = pt.byte_off[c][p + 1];
      RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(src->name(), d.type, 0, m, b1 - b0, d.validity != nullptr, stream, &col));
      CYL_RETURN_IF_FAILED(cyl_concat_offsets(d.offsets + r0, m, nullptr, 0, col->mutable_offsets(), stream));
      CYL_RETURN_IF_FAILED(cyl_memcpy_d2d(col->mutable_data(), static_cast<const char *>(d.data) + b0, b1 - b0, stream));
    } else {
      const int64_t w = d.byte_width;
      RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(src->name(), d.type, (int32_t)w, m, m * w, d.validity != nullptr, stream, &col));
      CYL_RETURN_IF_FAILED(cyl_memcpy_d2d(col->mutable_data(), static_cast<const char *>(d.data) + r0 * w, m * w, stream));
    }
    if (d.validity)  // re-align the bit range [r0, r1) to bit 0 through the byte form
      CYL_RETURN_IF_FAILED(cyl_pack_validity(validity_bytes[c]->as<uint8_t>() + r0, m, col->mutable_validity(), stream));
    cols.push_back(col);
  }
  *out = std::make_shared<Table>(std::move(cols), ctx);
  return Status::OK();
}

}  // namespace

// ---------------------------------------------------------------------------------------------- partition API
Status MapToHashPartitions(const std::shared_ptr<Table> &table, const std::vector<int32_t> &hash_column_idx, uint32_t num_partitions,
                           uint32_t **target_partitions, uint64_t **partition_hist) {
  std::vector<int> cols(hash_column_idx.begin(), hash_column_idx.end());
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*table, cols, "MapToHashPartitions"));
  void *stream = table->GetContext()->GetStream();
  void *pid = nullptr, *hist = nullptr;
  CYL_RETURN_IF_FAILED(cyl_malloc(&pid, (table->Rows() + 1) * 4));
  CYL_RETURN_IF_FAILED(cyl_malloc(&hist, num_partitions * 8));
  const auto kd = detail::Descs(*table, cols);
  int rc = cyl_hash_partition(kd.data(), (int32_t)kd.size(), table->Rows(), num_partitions, static_cast<uint32_t *>(pid),
                              static_cast<uint64_t *>(hist), stream);
  if (rc != CYL_OK) {
    cyl_free(pid);
    cyl_free(hist);
    return detail::FromCode(rc);
  }
  *target_partitions = static_cast<uint32_t *>(pid);
  *partition_hist = static_cast<uint64_t *>(hist);
  return Status::OK();
}

Status MapToHashPartitions(const std::shared_ptr<Table> &table, int32_t hash_column_idx, uint32_t num_partitions, uint32_t **target_partitions,
                           uint64_t **partition_hist) {
  return MapToHashPartitions(table, std::vector<int32_t>{hash_column_idx}, num_partitions, target_partitions, partition_hist);
}

// Split by given partition ids: the partition-ordered copy of the table and the partition boundaries (host). Row order inside a
// partition = input order.
static Status PartitionTableByPid(const std::shared_ptr<Table> &table, uint32_t P, const uint32_t *target_partitions, Partitioned *out) {
  Partitioned &pt = *out;
  auto ctx = table->GetContext();
  void *stream = ctx->GetStream();
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t n = table->Rows();
  DevBuf dest, part_offsets;
  RETURN_CYLON_STATUS_IF_FAILED(dest.alloc((size_t)(n + 1) * 8, stream));
  RETURN_CYLON_STATUS_IF_FAILED(part_offsets.alloc((P + 1) * 8, stream));
  CYL_RETURN_IF_FAILED(cyl_scatter_map(target_partitions, n, P, dest.as<int64_t>(), part_offsets.as<int64_t>(), stream));
  pt.cols.clear();
  pt.byte_off.assign((size_t)table->Columns(), {});
  for (const auto &c : table->GetColumns()) {
    const cyl_column_t &src = c->desc();
    std::shared_ptr<Column> d;
    RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(c->name(), c->type(), c->byte_width(), n, c->data_bytes(), c->has_validity(), stream, &d));
    if (c->is_var_width()) {
      CYL_RETURN_IF_FAILED(cyl_scatter_varwidth(src.offsets, static_cast<const uint8_t *>(src.data), dest.as<int64_t>(), n, d->mutable_offsets(),
                                                static_cast<uint8_t *>(d->mutable_data()), stream));
    } else {
      CYL_RETURN_IF_FAILED(cyl_scatter_fixed(src.data, src.byte_width, dest.as<int64_t>(), n, d->mutable_data(), stream));
    }
    if (src.validity) CYL_RETURN_IF_FAILED(cyl_scatter_validity(src.validity, dest.as<int64_t>(), n, d->mutable_validity(), stream));
    pt.cols.push_back(d);
  }
  pt.row_off.assign(P + 1, 0);
  CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(pt.row_off.data(), part_offsets.p, (P + 1) * 8, cudaMemcpyDeviceToHost, s));
  CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(s));
  return BoundaryByteOffsets(pt.cols, part_offsets.as<int64_t>(), P, stream, &pt.byte_off);
}

Status Split(const std::shared_ptr<Table> &table, uint32_t P, const uint32_t *target_partitions, std::vector<std::shared_ptr<Table>> &output) {
  auto ctx = table->GetContext();
  void *stream = ctx->GetStream();
  Partitioned pt;
  RETURN_CYLON_STATUS_IF_FAILED(PartitionTableByPid(table, P, target_partitions, &pt));
  output.clear();
  std::vector<std::unique_ptr<DevBuf>> vbytes;
  RETURN_CYLON_STATUS_IF_FAILED(UnpackValidity(pt, stream, &vbytes));
  for (uint32_t p = 0; p < P; p++) {
    std::shared_ptr<Table> t;
    RETURN_CYLON_STATUS_IF_FAILED(SliceRows(pt, vbytes, ctx, p, &t));
    output.push_back(t);
  }
  return Status::OK();
}

// ---- the reference's host-vector signatures (partition/partition.hpp:32-42,76-90; partition.cpp:66-217) over the device forms
Status MapToHashPartitions(const std::shared_ptr<Table> &table, const std::vector<int32_t> &hash_column_idx, uint32_t num_partitions,
                           std::vector<uint32_t> &target_partitions, std::vector<uint32_t> &partition_histogram) {
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  uint32_t *pid = nullptr;
  uint64_t *hist = nullptr;
  RETURN_CYLON_STATUS_IF_FAILED(MapToHashPartitions(table, hash_column_idx, num_partitions, &pid, &hist));
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
  if (e != cudaSuccess) return Status(Code::ExecutionError, std::string("MapToHashPartitions: ") + cudaGetErrorString(e));
  partition_histogram.assign(num_partitions, 0);
  for (uint32_t p = 0; p < num_partitions; p++) partition_histogram[p] = (uint32_t)h64[p];  // the reference's counters are 32-bit
  return Status::OK();
}

Status MapToHashPartitions(const std::shared_ptr<Table> &table, int32_t hash_column_idx, uint32_t num_partitions,
                           std::vector<uint32_t> &target_partitions, std::vector<uint32_t> &partition_histogram) {
  return MapToHashPartitions(table, std::vector<int32_t>{hash_column_idx}, num_partitions, target_partitions, partition_histogram);
}

Status Split(const std::shared_ptr<Table> &table, uint32_t num_partitions, const std::vector<uint32_t> &target_partitions,
             std::vector<std::shared_ptr<Table>> &output) {
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  if ((int64_t)target_partitions.size() != table->Rows()) return Status(Code::Invalid, "Split: target_partitions has a different length than the table");
  for (uint32_t p : target_partitions)
    if (p >= num_partitions) return Status(Code::IndexError, "Split: target partition out of range");
  void *stream = table->GetContext()->GetStream();
  DevBuf pid;
  RETURN_CYLON_STATUS_IF_FAILED(pid.alloc((target_partitions.size() + 1) * 4, stream));
  if (!target_partitions.empty())
    CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(pid.p, target_partitions.data(), target_partitions.size() * 4, cudaMemcpyHostToDevice,
                                          static_cast<cudaStream_t>(stream)));
  Status st = Split(table, num_partitions, pid.as<uint32_t>(), output);
  CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));  // the host vector may go away now
  return st;
}

Status Split(const std::shared_ptr<Table> &table, uint32_t num_partitions, const std::vector<uint32_t> &target_partitions,
             const std::vector<uint32_t> &partition_hist_ptr, std::vector<std::shared_ptr<Table>> &output) {
  if (partition_hist_ptr.size() != num_partitions) return Status(Code::Invalid, "Split: partition histogram length != num_partitions");
  RETURN_CYLON_STATUS_IF_FAILED(Split(table, num_partitions, target_partitions, output));
  for (uint32_t p = 0; p < num_partitions; p++)
    if (output[p]->Rows() != (int64_t)partition_hist_ptr[p]) return Status(Code::Invalid, "Split: partition histogram does not match target_partitions");
  return Status::OK();
}

Status PartitionByHashing(const std::shared_ptr<Table> &table, const std::vector<int32_t> &hash_cols, uint32_t num_partitions,
                          std::vector<std::shared_ptr<Table>> &partitions) {  // partition.cpp:200-217: MapToHashPartitions + Split
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  if (num_partitions < 1) return Status(Code::Invalid, "num_partitions must be positive");
  std::shared_ptr<Table> t = table;
  std::unordered_map<int, std::shared_ptr<Table>> parts;
  RETURN_CYLON_STATUS_IF_FAILED(HashPartition(t, std::vector<int>(hash_cols.begin(), hash_cols.end()), (int)num_partitions, &parts));
  partitions.clear();
  for (uint32_t p = 0; p < num_partitions; p++) partitions.push_back(parts[(int)p]);
  return Status::OK();
}

Status HashPartition(std::shared_ptr<Table> &table, const std::vector<int> &hash_columns, int no_of_partitions,
                     std::unordered_map<int, std::shared_ptr<Table>> *out) {  // table.cpp:384-426
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  if (no_of_partitions < 1) return Status(Code::Invalid, "no_of_partitions must be positive");
  Partitioned pt;
  RETURN_CYLON_STATUS_IF_FAILED(PartitionTable(table, hash_columns, (uint32_t)no_of_partitions, &pt));
  std::vector<std::unique_ptr<DevBuf>> vbytes;
  RETURN_CYLON_STATUS_IF_FAILED(UnpackValidity(pt, table->GetContext()->GetStream(), &vbytes));
  for (int p = 0; p < no_of_partitions; p++) {
    std::shared_ptr<Table> t;
    RETURN_CYLON_STATUS_IF_FAILED(SliceRows(pt, vbytes, table->GetContext(), (uint32_t)p, &t));
    out->insert({p, t});
  }
  return Status::OK();
}

// ---------------------------------------------------------------------------------------------- shuffle
}  // namespace cylon

// Pure host arithmetic of the all-to-all (no GPU, no communicator): counts[src * world + dst] = units (rows or bytes)
// that rank `src` sends to rank `dst`, as every rank knows after the count exchange. For `rank` it yields where each
// peer's slice starts in its partition-ordered send buffer and in its receive buffer (chunks ordered by source rank).
extern "C" int cyl_shuffle_plan(const int64_t *counts, int32_t world, int32_t rank, int64_t *send_off, int64_t *send_cnt, int64_t *recv_off,
                                int64_t *recv_cnt, int64_t *total_recv) {
  if (!counts || world < 1 || rank < 0 || rank >= world) return CYL_Invalid;
  int64_t so = 0, ro = 0;
  for (int p = 0; p < world; p++) {
    const int64_t s = counts[(int64_t)rank * world + p], r = counts[(int64_t)p * world + rank];
    if (s < 0 || r < 0) return CYL_Invalid;
    send_off[p] = so;
    send_cnt[p] = s;
    so += s;
    recv_off[p] = ro;
    recv_cnt[p] = r;
    ro += r;
  }
  if (total_recv) *total_recv = ro;
  return CYL_OK;
}

namespace cylon {

// Every rank holds the full counts matrix after the count exchange, so a capacity verdict can be reached for ALL
// receivers and returned by all ranks together; a rank-local early return between collectives would leave the others
// waiting in the next one.
static Status CheckReceiveTotals(const std::vector<int64_t> &counts, int P) {
  for (int dst = 0; dst < P; dst++) {
    int64_t rows = 0;
    for (int sr = 0; sr < P; sr++) rows += counts[(size_t)sr * P + dst];
    if (rows > 0xFFFFFFFEll) return Status(Code::CapacityError, "shuffle would put more than 2^32-2 rows on rank " + std::to_string(dst));
  }
  return Status::OK();
}

// Fused partition + shuffle over peer memory, for tables whose columns are all fixed width (1/2/4/8/16 bytes) and carry
// no nulls on any rank: histogram pass -> count exchange -> every rank allocates its output columns and maps the
// others' through CUDA IPC -> ONE scatter kernel per rank writes each row straight to its final position, locally or
// in a peer GPU's output column over NVLink. No send buffer, no NCCL payload traffic, no receive-side copy; the two
// StreamBarriers order the remote stores against the owners' streams. *handled == false: nothing was done and the
// caller must take the general path (the decision is the same on every rank).
static Status ShufflePeerScatter(std::shared_ptr<CylonContext> &ctx, std::shared_ptr<Table> &table, const std::vector<int> &hash_cols,
                                 std::shared_ptr<Table> &out, bool *handled) {
  *handled = false;
  if (ctx->GetConfig("cylon.shuffle.peer", "1") == "0") return Status::OK();
  for (const auto &c : table->GetColumns()) {  // schema-only test: identical on every rank
    const int w = c->byte_width();
    if (c->is_var_width() || !(w == 1 || w == 2 || w == 4 || w == 8 || w == 16)) return Status::OK();
  }
  auto comm = ctx->GetCommunicator();
  const int P = ctx->GetWorldSize(), rank = ctx->GetRank(), ncols = table->Columns();
  void *stream = ctx->GetStream();
  detail::StageTrace trace("shuffle(peer)", stream);

  // One exchange carries this rank's P partition counts AND its "some column has nulls" flag: the decision to stay on this
  // path needs agreement before any rank commits to it, and the histogram it costs to find out is cheap (0.16 ms per
  // 100M rows) next to a second round trip through the host on every shuffle.
  const auto all = detail::AllDescs(*table);
  std::vector<int32_t> keys(hash_cols.begin(), hash_cols.end());
  cyl_partition_plan_t *plan = nullptr;
  std::vector<int64_t> mine((size_t)P + 1), gathered((size_t)(P + 1) * P), counts((size_t)P * P);
  for (const auto &c : table->GetColumns()) mine[(size_t)P] |= c->has_validity() ? 1 : 0;
  if (mine[(size_t)P] == 0)  // a table with nulls is not planned at all; it still has to take part in the exchange
    CYL_RETURN_IF_FAILED(cyl_partition_plan_create(all.data(), ncols, keys.data(), (int32_t)keys.size(), table->Rows(), (uint32_t)P, &plan,
                                                   mine.data(), stream));
  struct PlanGuard {
    cyl_partition_plan_t *p;
    ~PlanGuard() { cyl_partition_plan_free(p); }
  } guard{plan};
  trace.mark("histogram");
  RETURN_CYLON_STATUS_IF_FAILED(comm->AllGather(mine.data(), P + 1, gathered.data(), stream));
  for (int sr = 0; sr < P; sr++) {
    if (gathered[(size_t)sr * (P + 1) + P]) return Status::OK();  // somebody has nulls: everybody takes the general path
    for (int d = 0; d < P; d++) counts[(size_t)sr * P + d] = gathered[(size_t)sr * (P + 1) + d];
  }
  std::vector<int64_t> so(P), sc(P), ro(P), rc(P);
  int64_t total_rows = 0;
  CYL_RETURN_IF_FAILED(cyl_shuffle_plan(counts.data(), P, rank, so.data(), sc.data(), ro.data(), rc.data(), &total_rows));
  RETURN_CYLON_STATUS_IF_FAILED(CheckReceiveTotals(counts, P));  // same verdict on every rank: nobody is left waiting

  // my output columns = what the peers write into
  std::vector<std::shared_ptr<Column>> out_cols((size_t)ncols);
  std::vector<void *> my_bufs((size_t)ncols), peer_bufs((size_t)ncols * P);
  for (int c = 0; c < ncols; c++) {
    const auto &src = table->GetColumn(c);
    RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(src->name(), src->type(), src->byte_width(), total_rows, total_rows * (int64_t)src->byte_width(),
                                                   false, stream, &out_cols[c]));
    my_bufs[c] = out_cols[c]->mutable_data();
  }
  bool all_ok = false;
  RETURN_CYLON_STATUS_IF_FAILED(comm->MapPeerBuffers(my_bufs.data(), ncols, peer_bufs.data(), &all_ok, stream));
  if (!all_ok) return Status::OK();  // agreed on every rank: general path
  trace.mark("count exchange + peer mapping");

  // where my rows start in rank g's output: after the rows of lower ranks (chunks ordered by source rank)
  std::vector<int64_t> dst_base((size_t)P, 0);
  for (int g = 0; g < P; g++)
    for (int sr = 0; sr < rank; sr++) dst_base[g] += counts[(size_t)sr * P + g];
  RETURN_CYLON_STATUS_IF_FAILED(comm->StreamBarrier(stream));  // every owner's stream is past the last use of its buffers
  CYL_RETURN_IF_FAILED(cyl_partition_plan_scatter(plan, peer_bufs.data(), dst_base.data(), stream));
  RETURN_CYLON_STATUS_IF_FAILED(comm->StreamBarrier(stream));  // all remote stores have landed before anyone reads
  trace.mark("scatter to peers");
  if (!table->IsRetain()) table.reset();
  out = std::make_shared<Table>(std::move(out_cols), ctx);
  *handled = true;
  return Status::OK();
}

// given_pid (device, one target rank per row; may be null): route by these ids instead of hashing hash_cols -- the range
// partitioning of DistributedSort. Such tables take the general path.
static Status ShuffleByHashing(std::shared_ptr<CylonContext> &ctx, std::shared_ptr<Table> &table, const std::vector<int> &hash_cols,
                               std::shared_ptr<Table> &out, const uint32_t *given_pid = nullptr) {
  auto comm = ctx->GetCommunicator();
  const int P = ctx->GetWorldSize(), rank = ctx->GetRank();
  void *stream = ctx->GetStream();
  const int ncols = table->Columns();

  if (!given_pid) {
    bool handled = false;
    RETURN_CYLON_STATUS_IF_FAILED(ShufflePeerScatter(ctx, table, hash_cols, out, &handled));
    if (handled) return Status::OK();
  }
  detail::StageTrace trace("shuffle", stream);
  Partitioned pt;
  if (given_pid) RETURN_CYLON_STATUS_IF_FAILED(PartitionTableByPid(table, (uint32_t)P, given_pid, &pt));
  else RETURN_CYLON_STATUS_IF_FAILED(PartitionTable(table, hash_cols, (uint32_t)P, &pt));
  trace.mark("partition");
  std::vector<std::string> names = table->ColumnNames();
  std::vector<int> has_validity(ncols);
  for (int c = 0; c < ncols; c++) has_validity[c] = table->GetColumn(c)->has_validity() ? 1 : 0;
  if (!table->IsRetain()) table.reset();  // table.cpp:152-154

  // ---- count exchange: [rows to p]*P, then per var-width column [bytes to p]*P, then has-nulls per column
  std::vector<int> var_cols;
  for (int c = 0; c < ncols; c++)
    if (pt.cols[c]->is_var_width()) var_cols.push_back(c);
  const int nvec = P * (1 + (int)var_cols.size()) + ncols;
  std::vector<int64_t> mine((size_t)nvec), all((size_t)nvec * P);
  for (int p = 0; p < P; p++) mine[p] = pt.row_off[p + 1] - pt.row_off[p];
  for (size_t v = 0; v < var_cols.size(); v++)
    for (int p = 0; p < P; p++) mine[P * (1 + v) + p] = pt.byte_off[var_cols[v]][p + 1] - pt.byte_off[var_cols[v]][p];
  for (int c = 0; c < ncols; c++) mine[P * (1 + var_cols.size()) + c] = has_validity[c];
  RETURN_CYLON_STATUS_IF_FAILED(comm->AllGather(mine.data(), nvec, all.data(), stream));
  trace.mark("count exchange");
  auto at = [&](int src, int field) { return all[(size_t)src * nvec + field]; };

  // ---- receive layout: chunks ordered by source rank (cyl_shuffle_plan)
  std::vector<int64_t> row_matrix((size_t)P * P);
  for (int sr = 0; sr < P; sr++)
    for (int d = 0; d < P; d++) row_matrix[(size_t)sr * P + d] = at(sr, d);
  std::vector<int64_t> recv_rows(P), recv_row_off(P + 1, 0), plan_send_off(P), plan_send_cnt(P);
  int64_t total_rows = 0;
  CYL_RETURN_IF_FAILED(cyl_shuffle_plan(row_matrix.data(), P, rank, plan_send_off.data(), plan_send_cnt.data(), recv_row_off.data(),
                                        recv_rows.data(), &total_rows));
  recv_row_off[P] = total_rows;
  RETURN_CYLON_STATUS_IF_FAILED(CheckReceiveTotals(row_matrix, P));  // same verdict on every rank
  for (size_t v = 0; v < var_cols.size(); v++) {  // int32 offsets: evaluated for EVERY receiver, so all ranks return alike
    for (int dst = 0; dst < P; dst++) {
      int64_t bytes = 0;
      for (int sr = 0; sr < P; sr++) bytes += at(sr, P * (1 + (int)v) + dst);
      if (bytes > 0x7FFFFFFFll)
        return Status(Code::CapacityError, "shuffled variable-width column exceeds 2^31-1 bytes on rank " + std::to_string(dst));
    }
  }

  std::vector<std::shared_ptr<Column>> out_cols((size_t)ncols);
  std::vector<net::Communicator::Exchange> ex;
  std::vector<std::unique_ptr<DevBuf>> keep;  // staging that must outlive the exchange
  struct Post { int col; DevBuf *lens; DevBuf *vbytes; };
  std::vector<Post> posts;

  auto add_exchange = [&](const void *send, void *recv, const std::vector<int64_t> &soff, const std::vector<int64_t> &sbytes,
                          const std::vector<int64_t> &roff, const std::vector<int64_t> &rbytes) {
    net::Communicator::Exchange e;
    e.send = send;
    e.recv = recv;
    e.send_off = soff;
    e.send_bytes = sbytes;
    e.recv_off = roff;
    e.recv_bytes = rbytes;
    ex.push_back(std::move(e));
  };
  auto scaled = [&](const std::vector<int64_t> &v, int64_t w, size_t n) {
    std::vector<int64_t> o(n);
    for (size_t i = 0; i < n; i++) o[i] = v[i] * w;
    return o;
  };
  std::vector<int64_t> send_rows(P), send_row_off(pt.row_off.begin(), pt.row_off.begin() + P);
  for (int p = 0; p < P; p++) send_rows[p] = pt.row_off[p + 1] - pt.row_off[p];
  std::vector<int64_t> recv_row_off_p(recv_row_off.begin(), recv_row_off.begin() + P);

  int vidx = 0;
  for (int c = 0; c < ncols; c++) {
    const auto &src = pt.cols[c];
    bool any_validity = false;
    for (int sr = 0; sr < P; sr++) any_validity |= at(sr, P * (1 + (int)var_cols.size()) + c) != 0;
    Post post{c, nullptr, nullptr};
    if (src->is_var_width()) {
      std::vector<int64_t> sbytes(P), soff(P), rbytes(P), roff(P + 1, 0);
      for (int p = 0; p < P; p++) {
        sbytes[p] = mine[P * (1 + vidx) + p];
        soff[p] = pt.byte_off[c][p];
        rbytes[p] = at(p, P * (1 + vidx) + rank);
        roff[p + 1] = roff[p] + rbytes[p];
      }
      RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(names[c], src->type(), 0, total_rows, roff[P], any_validity, stream, &out_cols[c]));
      // lengths travel as int32 per row; offsets are rebuilt on the receiver
      keep.emplace_back(new DevBuf());
      DevBuf *slens = keep.back().get();
      keep.emplace_back(new DevBuf());
      DevBuf *rlens = keep.back().get();
      RETURN_CYLON_STATUS_IF_FAILED(slens->alloc((size_t)(src->length() + 1) * 4, stream));
      RETURN_CYLON_STATUS_IF_FAILED(rlens->alloc((size_t)(total_rows + 1) * 4, stream));
      CYL_RETURN_IF_FAILED(cyl_offsets_to_lengths(src->desc().offsets, src->length(), slens->as<int32_t>(), stream));
      add_exchange(slens->p, rlens->p, scaled(send_row_off, 4, P), scaled(send_rows, 4, P), scaled(recv_row_off_p, 4, P), scaled(recv_rows, 4, P));
      std::vector<int64_t> roff_p(roff.begin(), roff.begin() + P);
      add_exchange(src->desc().data, out_cols[c]->mutable_data(), soff, sbytes, roff_p, rbytes);
      post.lens = rlens;
      vidx++;
    } else {
      const int64_t w = src->byte_width();
      RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(names[c], src->type(), (int32_t)w, total_rows, total_rows * w, any_validity, stream, &out_cols[c]));
      add_exchange(src->desc().data, out_cols[c]->mutable_data(), scaled(send_row_off, w, P), scaled(send_rows, w, P), scaled(recv_row_off_p, w, P),
                   scaled(recv_rows, w, P));
    }
    if (any_validity) {
      keep.emplace_back(new DevBuf());
      DevBuf *sv = keep.back().get();
      keep.emplace_back(new DevBuf());
      DevBuf *rv = keep.back().get();
      RETURN_CYLON_STATUS_IF_FAILED(sv->alloc((size_t)src->length() + 1, stream));
      RETURN_CYLON_STATUS_IF_FAILED(rv->alloc((size_t)total_rows + 1, stream));
      CYL_RETURN_IF_FAILED(cyl_unpack_validity(src->desc().validity, src->length(), sv->as<uint8_t>(), stream));  // null bitmap -> all ones
      add_exchange(sv->p, rv->p, send_row_off, send_rows, recv_row_off_p, recv_rows);
      post.vbytes = rv;
    }
    posts.push_back(post);
  }

  trace.mark("allocate outputs");
  RETURN_CYLON_STATUS_IF_FAILED(comm->AllToAllV(ex, stream));
  trace.mark("all-to-all");

  for (const auto &post : posts) {
    if (post.lens)
      CYL_RETURN_IF_FAILED(cyl_lengths_to_offsets(post.lens->as<int32_t>(), total_rows, out_cols[post.col]->mutable_offsets(), nullptr, stream));
    if (post.vbytes) CYL_RETURN_IF_FAILED(cyl_pack_validity(post.vbytes->as<uint8_t>(), total_rows, out_cols[post.col]->mutable_validity(), stream));
  }
  out = std::make_shared<Table>(std::move(out_cols), ctx);
  return Status::OK();
}

namespace detail {
Status ShuffleByPartitionIds(std::shared_ptr<CylonContext> &ctx, std::shared_ptr<Table> &table, const uint32_t *target_ranks, std::shared_ptr<Table> &out) {
  return ShuffleByHashing(ctx, table, {}, out, target_ranks);
}
}  // namespace detail

Status Shuffle(std::shared_ptr<Table> &table, const std::vector<int> &hash_columns, std::shared_ptr<Table> &output) {  // table.cpp:910-921
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  auto ctx = table->GetContext();
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*table, hash_columns, "Shuffle"));
  if (ctx->GetWorldSize() == 1) {  // one partition routed to ourselves: nothing to move
    output = table;
    return Status::OK();
  }
  return ShuffleByHashing(ctx, table, hash_columns, output);
}

static Status ShuffleTwo(std::shared_ptr<CylonContext> &ctx, std::shared_ptr<Table> &left, const std::vector<int> &lcols,
                         std::shared_ptr<Table> &right, const std::vector<int> &rcols, std::shared_ptr<Table> &lout, std::shared_ptr<Table> &rout) {
  RETURN_CYLON_STATUS_IF_FAILED(ShuffleByHashing(ctx, left, lcols, lout));  // table.cpp:167-173
  return ShuffleByHashing(ctx, right, rcols, rout);
}

// INNER join of all-8-byte tables on one key (the benchmark shape): the left table is shuffled first and radix-partitioned
// into the join's bins on the context's SECOND stream while the right table is still being shuffled on the first -- its
// histogram and the host round trips of its count exchange leave the GPU idle otherwise, and the NVLink-bound scatter and
// the HBM-bound partition passes want different resources. The reference runs the two shuffles and the join strictly one
// after the other (table.cpp:159-179,482-502). *done == false: the shape does not qualify, nothing was touched.
// Opt-in ("cylon.join.overlap" = "1"): measured on 2 and 4 B200s it moves the step by under 1 % (8.66 -> 8.58 ms, 10.39 ->
// 10.32 ms) -- the kernels of a step already add up to its length, and all of them are bandwidth-bound on the same HBM.
static Status OverlappedInnerJoin(std::shared_ptr<CylonContext> &ctx, std::shared_ptr<Table> &left, std::shared_ptr<Table> &right,
                                  const join::config::JoinConfig &cfg, std::shared_ptr<Table> &out, bool *done) {
  *done = false;
  const auto &lidx = cfg.GetLeftColumnIdx(), &ridx = cfg.GetRightColumnIdx();
  if (cfg.GetType() != join::config::INNER || lidx.size() != 1 || ridx.size() != 1) return Status::OK();
  if (ctx->GetConfig("cylon.join.fused", "1") == "0" || ctx->GetConfig("cylon.join.overlap", "0") != "1") return Status::OK();
  {
    const auto la = detail::AllDescs(*left), ra = detail::AllDescs(*right);
    if (!cyl_hash_join_fused_supported(la.data(), (int32_t)la.size(), lidx[0], ra.data(), (int32_t)ra.size(), ridx[0], CYL_JOIN_INNER)) return Status::OK();
    for (const auto &t : {left, right})  // the two-step entry points take the all-8-byte shape only
      for (const auto &c : t->GetColumns())
        if (c->byte_width() != 8) return Status::OK();
  }
  cudaStream_t s = static_cast<cudaStream_t>(ctx->GetStream());
  cudaStream_t aux = static_cast<cudaStream_t>(ctx->GetAuxStream());
  if (!aux) return Status::OK();
  *done = true;

  struct Events {
    cudaEvent_t e[3] = {nullptr, nullptr, nullptr};
    ~Events() {
      for (auto ev : e)
        if (ev) cudaEventDestroy(ev);
    }
  } ev;
  for (auto &e : ev.e) CUDA_RETURN_IF_FAILED(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  struct Sides {
    cyl_join_side_t *l = nullptr, *r = nullptr;
    ~Sides() {
      cyl_hash_join_fused_side_free(l);
      cyl_hash_join_fused_side_free(r);
    }
  } sides;

  std::shared_ptr<Table> l, r;
  const std::string lprefix = cfg.GetLeftTableSuffix(), rprefix = cfg.GetRightTableSuffix();
  RETURN_CYLON_STATUS_IF_FAILED(ShuffleByHashing(ctx, left, lidx, l));
  CUDA_RETURN_IF_FAILED(cudaEventRecord(ev.e[0], s));
  CUDA_RETURN_IF_FAILED(cudaStreamWaitEvent(aux, ev.e[0], 0));
  const auto lall = detail::AllDescs(*l);
  CYL_RETURN_IF_FAILED(cyl_hash_join_fused_partition(lall.data(), (int32_t)lall.size(), lidx[0], l->Rows(), nullptr, l->Rows(), &sides.l, aux));
  CUDA_RETURN_IF_FAILED(cudaEventRecord(ev.e[1], aux));
  RETURN_CYLON_STATUS_IF_FAILED(ShuffleByHashing(ctx, right, ridx, r));
  const auto rall = detail::AllDescs(*r);
  CYL_RETURN_IF_FAILED(cyl_hash_join_fused_partition(rall.data(), (int32_t)rall.size(), ridx[0], r->Rows(), sides.l, 0, &sides.r, s));
  CUDA_RETURN_IF_FAILED(cudaStreamWaitEvent(s, ev.e[1], 0));

  const int64_t nl = l->Rows(), nr = r->Rows();
  int64_t capacity = std::max(nl, nr) + std::max(nl, nr) / 8 + 1024;
  Status result(Code::ExecutionError, "fused join did not fit its exact capacity");
  for (int attempt = 0; attempt < 2; attempt++) {
    std::vector<std::shared_ptr<Column>> cols;
    std::vector<void *> outs;
    for (int side = 0; side < 2; side++) {
      const Table &t = side == 0 ? *l : *r;
      for (const auto &src : t.GetColumns()) {
        std::shared_ptr<Column> col;
        RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate((side == 0 ? lprefix : rprefix) + src->name(), src->type(), 8, capacity, capacity * 8, false,
                                                       ctx->GetStream(), &col));
        outs.push_back(col->mutable_data());
        cols.push_back(col);
      }
    }
    int64_t m = 0;
    CYL_RETURN_IF_FAILED(cyl_hash_join_fused_probe(sides.l, sides.r, capacity, outs.data(), &m, ctx->GetStream()));
    if (m <= capacity) {
      for (auto &c : cols) c->truncate(m);
      out = std::make_shared<Table>(std::move(cols), ctx);
      result = Status::OK();
      break;
    }
    capacity = m;
  }
  // the left side's scratch goes back to the second stream's free lists: later work there must come after this probe
  CUDA_RETURN_IF_FAILED(cudaEventRecord(ev.e[2], s));
  CUDA_RETURN_IF_FAILED(cudaStreamWaitEvent(aux, ev.e[2], 0));
  return result;
}

Status DistributedJoin(std::shared_ptr<Table> &left, std::shared_ptr<Table> &right, const join::config::JoinConfig &cfg,
                       std::shared_ptr<Table> &out) {
  if (left == nullptr) return Status(Code::KeyError, "Couldn't find the left table");
  if (right == nullptr) return Status(Code::KeyError, "Couldn't find the right table");
  auto ctx = left->GetContext();
  if (ctx->GetWorldSize() == 1) return Join(left, right, cfg, out);  // table.cpp:487
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*left, cfg.GetLeftColumnIdx(), "DistributedJoin(left)"));
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*right, cfg.GetRightColumnIdx(), "DistributedJoin(right)"));
  std::shared_ptr<Table> l, r;
  detail::StageTrace trace("DistributedJoin", ctx->GetStream());
  {
    bool done = false;
    RETURN_CYLON_STATUS_IF_FAILED(OverlappedInnerJoin(ctx, left, right, cfg, out, &done));
    if (done) return Status::OK();
  }
  RETURN_CYLON_STATUS_IF_FAILED(ShuffleTwo(ctx, left, cfg.GetLeftColumnIdx(), right, cfg.GetRightColumnIdx(), l, r));
  trace.mark("shuffle both tables");
  Status st = Join(l, r, cfg, out);
  trace.mark("local join");
  return st;
}

typedef Status (*LocalSetOperation)(std::shared_ptr<Table> &, std::shared_ptr<Table> &, std::shared_ptr<Table> &);

static Status DoDistributedSetOperation(LocalSetOperation op, std::shared_ptr<Table> &left, std::shared_ptr<Table> &right,
                                        std::shared_ptr<Table> &out) {  // table.cpp:727-767
  if (left == nullptr || right == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  RETURN_CYLON_STATUS_IF_FAILED(detail::VerifyTableSchema(*left, *right));
  auto ctx = left->GetContext();
  if (ctx->GetWorldSize() < 2) return op(left, right, out);
  std::vector<int> all_cols;
  for (int c = 0; c < left->Columns(); c++) all_cols.push_back(c);
  std::shared_ptr<Table> l, r;
  RETURN_CYLON_STATUS_IF_FAILED(ShuffleTwo(ctx, left, all_cols, right, all_cols, l, r));
  return op(l, r, out);
}

Status DistributedUnion(std::shared_ptr<Table> &l, std::shared_ptr<Table> &r, std::shared_ptr<Table> &out) {
  return DoDistributedSetOperation(&Union, l, r, out);
}
Status DistributedSubtract(std::shared_ptr<Table> &l, std::shared_ptr<Table> &r, std::shared_ptr<Table> &out) {
  return DoDistributedSetOperation(&Subtract, l, r, out);
}
Status DistributedIntersect(std::shared_ptr<Table> &l, std::shared_ptr<Table> &r, std::shared_ptr<Table> &out) {
  return DoDistributedSetOperation(&Intersect, l, r, out);
}

Status DistributedUnique(std::shared_ptr<Table> &in, const std::vector<int> &cols, std::shared_ptr<Table> &out) {  // table.cpp:984-999
  if (in == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  auto ctx = in->GetContext();
  if (ctx->GetWorldSize() == 1) return Unique(in, cols, out);
  std::shared_ptr<Table> shuffled;
  RETURN_CYLON_STATUS_IF_FAILED(Shuffle(in, cols, shuffled));
  return Unique(shuffled, cols, out);
}

}  // namespace cylon
