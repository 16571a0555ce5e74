// extern "C" doors onto the reference's OWN compiled code (hash partition kernels, row hashing and the single-key
// hash join), built from the sources where they lie under /root/reference against pyarrow's Arrow C++ (24.x).
// CHECKER ONLY: used by tests/ to pin the oracle and by bench.py as the "reference" CPU baseline. Nothing here is
// copied from the reference; this file only wraps raw buffers into arrow::Array objects and calls it.
#include <arrow/api.h>
#include <arrow/compute/api.h>

#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "arrow/arrow_partition_kernels.hpp"
#include "join/hash_join.hpp"
#include "join/join.hpp"
#include "arrow/arrow_types.hpp"
#include "compute/aggregates.hpp"
#include "join/join_config.hpp"
#include "net/mpi/mpi_operations.hpp"
#include "util/arrow_utils.hpp"

// ---- link-time stand-ins for reference functions that are reachable but never called by the checker ----------
namespace cylon {
namespace mpi {
cylon::Status AllReduce(const void *, void *, int, const std::shared_ptr<DataType> &, cylon::net::ReduceOp) {
  return cylon::Status(cylon::Code::NotImplemented, "no MPI in the checker build");
}
}  // namespace mpi
namespace util {
// util/copy_arrray.cpp does not compile against Arrow 24 (NumericBuilder<BooleanType> no longer exists), so the
// checker supplies the one function the join code needs from it, with the documented semantics
// (util/copy_arrray.cpp:24-110: index -1 -> null, otherwise the value) expressed through arrow::compute::Take.
// The join algorithms themselves (hash_join.cpp, sort_join.cpp, join_utils.cpp) are the reference's own code.
arrow::Status copy_array_by_indices(const std::vector<int64_t> &indices, const std::shared_ptr<arrow::Array> &data_array,
                                    std::shared_ptr<arrow::Array> *copied_array, arrow::MemoryPool *memory_pool) {
  arrow::Int64Builder b(memory_pool);
  ARROW_RETURN_NOT_OK(b.Reserve((int64_t)indices.size()));
  for (int64_t v : indices) {
    if (v < 0) b.UnsafeAppendNull(); else b.UnsafeAppend(v);
  }
  std::shared_ptr<arrow::Array> idx;
  ARROW_RETURN_NOT_OK(b.Finish(&idx));
  ARROW_ASSIGN_OR_RAISE(auto taken, arrow::compute::Take(arrow::Datum(data_array), arrow::Datum(idx)));
  *copied_array = taken.make_array();
  return arrow::Status::OK();
}
}  // namespace util
arrow::MemoryPool *ToArrowPool(const std::shared_ptr<cylon::CylonContext> &) { return arrow::default_memory_pool(); }
// referenced only by the range-partition kernel (DistributedSort), which the checker never runs
int CylonContext::GetWorldSize() { return 1; }
namespace tarrow {
std::shared_ptr<cylon::DataType> ToCylonType(const std::shared_ptr<arrow::DataType> &) { return nullptr; }
}  // namespace tarrow
namespace compute {
cylon::Status MinMax(std::shared_ptr<cylon::CylonContext> &, const arrow::Datum &, const std::shared_ptr<cylon::DataType> &,
                     std::shared_ptr<cylon::compute::Result> &) {
  return cylon::Status(cylon::Code::NotImplemented, "not part of the checker build");
}
}  // namespace compute
}  // namespace cylon

namespace {

std::shared_ptr<arrow::DataType> type_of(int32_t t) {
  switch (t) {
    case 2: return arrow::uint8();
    case 3: return arrow::int8();
    case 4: return arrow::uint16();
    case 5: return arrow::int16();
    case 6: return arrow::uint32();
    case 7: return arrow::int32();
    case 8: return arrow::uint64();
    case 9: return arrow::int64();
    case 11: return arrow::float32();
    case 12: return arrow::float64();
    case 13: return arrow::utf8();
    case 14: return arrow::binary();
    default: return nullptr;
  }
}

// wrap (no copy) raw host buffers into an arrow::Array
std::shared_ptr<arrow::Array> wrap(int32_t type, int32_t width, int64_t n, const int32_t *offsets, const void *data) {
  auto dt = type_of(type);
  if (!dt) return nullptr;
  std::vector<std::shared_ptr<arrow::Buffer>> bufs;
  bufs.push_back(nullptr);
  if (type == 13 || type == 14) {
    bufs.push_back(arrow::Buffer::Wrap(offsets, (size_t)n + 1));
    bufs.push_back(std::make_shared<arrow::Buffer>(static_cast<const uint8_t *>(data), n > 0 ? (int64_t)offsets[n] : 0));
  } else {
    bufs.push_back(std::make_shared<arrow::Buffer>(static_cast<const uint8_t *>(data), n * (int64_t)width));
  }
  return arrow::MakeArray(arrow::ArrayData::Make(dt, n, bufs, 0));
}

}  // namespace

extern "C" {

// HashPartitionKernel::UpdateHash of the reference (arrow_partition_kernels.cpp), in place on h
int ref_update_hash(int32_t type, int32_t width, int64_t n, const int32_t *offsets, const void *data, uint32_t *h) {
  auto arr = wrap(type, width, n, offsets, data);
  if (!arr) return 42;
  auto kern = cylon::CreateHashPartitionKernel(arr->type());
  if (!kern) return 42;
  std::vector<uint32_t> hashes(h, h + n);
  auto st = kern->UpdateHash(std::make_shared<arrow::ChunkedArray>(arr), hashes);
  if (!st.is_ok()) return st.get_code();
  std::memcpy(h, hashes.data(), (size_t)n * 4);
  return 0;
}

// HashPartitionKernel::Partition of the reference: pid in/out (running hash in, partition id out), hist out
int ref_partition(int32_t type, int32_t width, int64_t n, const int32_t *offsets, const void *data, uint32_t num_partitions, uint32_t *pid,
                  uint32_t *hist) {
  auto arr = wrap(type, width, n, offsets, data);
  if (!arr) return 42;
  auto kern = cylon::CreateHashPartitionKernel(arr->type());
  if (!kern) return 42;
  std::vector<uint32_t> p(pid, pid + n), hgram(num_partitions, 0);
  auto st = kern->Partition(std::make_shared<arrow::ChunkedArray>(arr), num_partitions, p, hgram);
  if (!st.is_ok()) return st.get_code();
  std::memcpy(pid, p.data(), (size_t)n * 4);
  std::memcpy(hist, hgram.data(), (size_t)num_partitions * 4);
  return 0;
}

// cylon::join::ArrayIndexHashJoin of the reference (join/hash_join.cpp:188-307). Output vectors are malloc'ed.
int ref_array_index_hash_join(int32_t type, int32_t width, int64_t nl, const int32_t *loff, const void *ldata, int64_t nr,
                              const int32_t *roff, const void *rdata, int32_t join_type, int64_t **l_idx, int64_t **r_idx, int64_t *m) {
  auto l = wrap(type, width, nl, loff, ldata), r = wrap(type, width, nr, roff, rdata);
  if (!l || !r) return 42;
  std::vector<int64_t> li, ri;
  auto st = cylon::join::ArrayIndexHashJoin(l, r, static_cast<cylon::join::config::JoinType>(join_type), li, ri);
  if (!st.ok()) return (int)st.code();
  *m = (int64_t)li.size();
  *l_idx = static_cast<int64_t *>(malloc((li.size() + 1) * 8));
  *r_idx = static_cast<int64_t *>(malloc((ri.size() + 1) * 8));
  std::memcpy(*l_idx, li.data(), li.size() * 8);
  std::memcpy(*r_idx, ri.data(), ri.size() * 8);
  return 0;
}

void ref_free(void *p) { free(p); }

// ---- whole-table joins through cylon::join::JoinTables (join/join.cpp:80-100): HashJoin incl. multi-column keys
// (hash_join.cpp:92-186) and SortJoin (sort_join.cpp:36-721), result built by build_final_table (join_utils.cpp:126-181)
struct ref_table {
  std::shared_ptr<arrow::Table> t;
};

struct ref_col_in {
  int32_t type, width;
  int64_t n;
  const int32_t *offsets;
  const void *data;
};

// The reference's single-key SORT join sorts the key column IN PLACE, so tables handed to it own
// mutable copies of the caller's buffers.
static std::shared_ptr<arrow::Array> own(int32_t type, int32_t width, int64_t n, const int32_t *offsets, const void *data) {
  auto dt = type_of(type);
  if (!dt) return nullptr;
  auto copy = [](const void *p, int64_t bytes) -> std::shared_ptr<arrow::Buffer> {
    auto r = arrow::AllocateBuffer(bytes > 0 ? bytes : 8);
    if (!r.ok()) return nullptr;
    std::shared_ptr<arrow::Buffer> b = std::move(r).ValueOrDie();
    if (bytes > 0) std::memcpy(b->mutable_data(), p, (size_t)bytes);
    return b;
  };
  std::vector<std::shared_ptr<arrow::Buffer>> bufs;
  bufs.push_back(nullptr);
  if (type == 13 || type == 14) {
    bufs.push_back(copy(offsets, (n + 1) * 4));
    bufs.push_back(copy(data, n > 0 ? (int64_t)offsets[n] : 0));
  } else {
    bufs.push_back(copy(data, n * (int64_t)width));
  }
  return arrow::MakeArray(arrow::ArrayData::Make(dt, n, bufs, 0));
}

static std::shared_ptr<arrow::Table> make_table(const ref_col_in *cols, int32_t ncols) {
  std::vector<std::shared_ptr<arrow::Field>> fields;
  std::vector<std::shared_ptr<arrow::Array>> arrays;
  for (int c = 0; c < ncols; c++) {
    auto a = own(cols[c].type, cols[c].width, cols[c].n, cols[c].offsets, cols[c].data);
    if (!a) return nullptr;
    fields.push_back(arrow::field("c" + std::to_string(c), a->type()));
    arrays.push_back(a);
  }
  return arrow::Table::Make(arrow::schema(fields), arrays);
}

int ref_join_tables(const ref_col_in *lcols, int32_t nl_cols, const ref_col_in *rcols, int32_t nr_cols, const int32_t *lkeys,
                    const int32_t *rkeys, int32_t nkeys, int32_t join_type, int32_t algorithm, ref_table **out) {
  auto l = make_table(lcols, nl_cols), r = make_table(rcols, nr_cols);
  if (!l || !r) return 42;
  cylon::join::config::JoinConfig cfg(static_cast<cylon::join::config::JoinType>(join_type), std::vector<int>(lkeys, lkeys + nkeys),
                                      std::vector<int>(rkeys, rkeys + nkeys), static_cast<cylon::join::config::JoinAlgorithm>(algorithm), "l_",
                                      "r_");
  std::shared_ptr<arrow::Table> joined;
  auto st = cylon::join::JoinTables(l, r, cfg, &joined, arrow::default_memory_pool());
  if (!st.ok()) {
    fprintf(stderr, "ref_join_tables: %s\n", st.ToString().c_str());
    return (int)st.code();
  }
  auto comb = joined->CombineChunks();
  if (!comb.ok()) return 42;
  *out = new ref_table{comb.ValueOrDie()};
  return 0;
}

int64_t ref_table_rows(const ref_table *t) { return t->t->num_rows(); }
int32_t ref_table_cols(const ref_table *t) { return t->t->num_columns(); }
// bytes of the data buffer of column c (value bytes for var-width)
int64_t ref_table_col_bytes(const ref_table *t, int32_t c) {
  if (t->t->num_rows() == 0) return 0;
  auto a = t->t->column(c)->chunk(0);
  if (a->type_id() == arrow::Type::STRING || a->type_id() == arrow::Type::BINARY) {
    auto b = std::static_pointer_cast<arrow::BinaryArray>(a);
    return b->value_offset(b->length()) - b->value_offset(0);
  }
  return a->length() * (int64_t)a->type()->byte_width();
}
// copies column c out: data, int32 offsets (var-width, rebased to 0; may be NULL) and one validity byte per row
int ref_table_copy_col(const ref_table *t, int32_t c, void *data, int32_t *offsets, uint8_t *valid) {
  if (t->t->num_rows() == 0) return 0;
  auto a = t->t->column(c)->chunk(0);
  const int64_t n = a->length();
  for (int64_t i = 0; i < n; i++) valid[i] = a->IsValid(i) ? 1 : 0;
  if (a->type_id() == arrow::Type::STRING || a->type_id() == arrow::Type::BINARY) {
    auto b = std::static_pointer_cast<arrow::BinaryArray>(a);
    const int32_t base = b->value_offset(0);
    for (int64_t i = 0; i <= n; i++) offsets[i] = b->value_offset(i) - base;
    std::memcpy(data, b->value_data()->data() + base, (size_t)(b->value_offset(n) - base));
  } else {
    const int w = a->type()->byte_width();
    std::memcpy(data, a->data()->buffers[1]->data() + a->offset() * w, (size_t)n * w);
  }
  return 0;
}
void ref_table_free(ref_table *t) { delete t; }

}  // extern "C"
