// extern "C" doors onto the reference's OWN compiled code (hash partition kernels, row hashing and the single-key
// hash join), built from the sources where they lie under /root/reference against pyarrow's Arrow C++ (24.x).
// CHECKER ONLY: used by tests/ to pin the oracle and by bench.py as the "reference" CPU baseline. Nothing here is
// copied from the reference; this file only wraps raw buffers into arrow::Array objects and calls it.
#include <arrow/api.h>
#include <arrow/compute/api.h>

#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <vector>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <thread>

#include "arrow/arrow_partition_kernels.hpp"
#include "ctx/cylon_context.hpp"
#include "groupby/groupby.hpp"
#include "groupby/pipeline_groupby.hpp"
#include "io/arrow_io.hpp"
#include "net/channel.hpp"
#include "net/communicator.hpp"
#include "table.hpp"
#include "join/hash_join.hpp"
#include "join/join.hpp"
#include "arrow/arrow_types.hpp"
#include "compute/aggregates.hpp"
#include "join/join_config.hpp"
#include "net/mpi/mpi_operations.hpp"
#include "partition/partition.hpp"
#include "util/arrow_utils.hpp"

// ---- link-time stand-ins for reference functions that are reachable but never called by the checker ----------
namespace cylon {
arrow::MemoryPool *ToArrowPool(const std::shared_ptr<cylon::CylonContext> &) { return arrow::default_memory_pool(); }
// The context without MPI (ctx/cylon_context.cpp pulls in the MPI communicator): same observable behaviour for the
// members the table operators use.
CylonContext::CylonContext(bool distributed) { this->is_distributed = distributed; }
std::shared_ptr<CylonContext> CylonContext::Init() { return std::make_shared<CylonContext>(false); }
int CylonContext::GetRank() { return is_distributed ? communicator->GetRank() : 0; }
int CylonContext::GetWorldSize() { return is_distributed ? communicator->GetWorldSize() : 1; }
std::vector<int> CylonContext::GetNeighbours(bool include_self) {
  std::vector<int> out;
  const int w = GetWorldSize(), me = GetRank();
  for (int r = 0; r < w; r++)
    if (include_self || r != me) out.push_back(r);
  return out;
}
int32_t CylonContext::GetNextSequence() { return sequence_no++; }
std::shared_ptr<net::Communicator> CylonContext::GetCommunicator() const { return communicator; }
void CylonContext::setCommunicator(const std::shared_ptr<cylon::net::Communicator> &c) { communicator = c; }
void CylonContext::setDistributed(bool d) { is_distributed = d; }
bool CylonContext::IsDistributed() const { return is_distributed; }
Status PipelineGroupBy(std::shared_ptr<Table> &, int32_t, const std::vector<std::pair<int32_t, compute::AggregationOpId>> &,
                       std::shared_ptr<Table> &) {
  return Status(Code::NotImplemented, "pipeline group-by is not part of the checker build");
}
namespace io {
arrow::Result<std::shared_ptr<arrow::Table>> read_csv(std::shared_ptr<cylon::CylonContext> &, const std::string &,
                                                      cylon::io::config::CSVReadOptions) {
  return arrow::Status::NotImplemented("no CSV reader in the checker build");
}
}  // namespace io
namespace util {
std::string generate_uuid_v4() {
  static std::atomic<uint64_t> n{0};
  return "checker-" + std::to_string(n++);
}
}  // namespace util
namespace tarrow {
std::shared_ptr<cylon::DataType> ToCylonType(const std::shared_ptr<arrow::DataType> &) { return nullptr; }
// arrow/arrow_types.cpp does not compile against Arrow 24 (DecimalType constructor). The table code only asks it whether
// a schema is made of types Cylon handles; the checker builds tables from the fixed list in type_of() below.
bool validateArrowTableTypes(const std::shared_ptr<arrow::Table> &) { return true; }
std::shared_ptr<arrow::DataType> convertToArrowType(const std::shared_ptr<DataType> &, int32_t, int32_t, int32_t) { return nullptr; }
}  // namespace tarrow
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
  const uint8_t *validity;  // Arrow validity bitmap (LSB first) or NULL = no nulls
};

// The reference's single-key SORT join sorts the key column IN PLACE, so tables handed to it own
// mutable copies of the caller's buffers.
static std::shared_ptr<arrow::Array> own(int32_t type, int32_t width, int64_t n, const int32_t *offsets, const void *data,
                                         const uint8_t *validity = nullptr) {
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
  bufs.push_back(validity ? copy(validity, (n + 7) / 8) : nullptr);
  if (type == 13 || type == 14) {
    bufs.push_back(copy(offsets, (n + 1) * 4));
    bufs.push_back(copy(data, n > 0 ? (int64_t)offsets[n] : 0));
  } else {
    bufs.push_back(copy(data, n * (int64_t)width));
  }
  return arrow::MakeArray(arrow::ArrayData::Make(dt, n, bufs, validity ? arrow::kUnknownNullCount : 0));
}

static std::shared_ptr<arrow::Table> make_table(const ref_col_in *cols, int32_t ncols) {
  std::vector<std::shared_ptr<arrow::Field>> fields;
  std::vector<std::shared_ptr<arrow::Array>> arrays;
  for (int c = 0; c < ncols; c++) {
    auto a = own(cols[c].type, cols[c].width, cols[c].n, cols[c].offsets, cols[c].data, cols[c].validity);
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
// name of column c (valid until the table is freed): the reference's prefixes ("l_", "sum_", "sum_sum_" ...)
const char *ref_table_col_name(const ref_table *t, int32_t c) { return t->t->field(c)->name().c_str(); }
int32_t ref_table_col_type(const ref_table *t, int32_t c) { return (int32_t)t->t->field(c)->type()->id(); }  // arrow::Type::type

}  // extern "C"

// ---- ranks as threads: an in-process Communicator/Channel so that the reference's Distributed* operators run here ------
// What MPIChannel shows its callbacks (net/mpi/mpi_channel.cpp:70-205) is kept: per message receivedHeader(source, 0,
// new int[h], h) then receivedData(source, buffer from the Allocator, length) -- also for 0-length buffers -- and for
// the end marker receivedHeader(source, 1, nullptr, 0); sends complete in the order they were accepted, the finish
// marker after all of them. Messages between two ranks on one edge stay in order.
namespace {

struct MemMsg {
  bool fin = false;
  std::vector<int> header;
  std::vector<uint8_t> data;
};

struct MemHub {
  std::mutex mu;
  std::map<std::tuple<int, int, int>, std::deque<MemMsg>> queues;  // (edge, source, target)
  std::condition_variable cv;
  int world = 1, waiting = 0;
  uint64_t generation = 0;
  void barrier() {
    std::unique_lock<std::mutex> lk(mu);
    const uint64_t gen = generation;
    if (++waiting == world) {
      waiting = 0;
      generation++;
      cv.notify_all();
    } else {
      cv.wait(lk, [&] { return generation != gen; });
    }
  }
  // every rank deposits n bytes and gets everybody's (what MPI_Allreduce / MPI_Allgather carry in the reference)
  std::vector<std::vector<uint8_t>> slots;
  std::vector<std::vector<uint8_t>> exchange(int rank, const void *p, size_t n) {
    {
      std::lock_guard<std::mutex> lk(mu);
      if ((int)slots.size() != world) slots.resize((size_t)world);
      const uint8_t *b = static_cast<const uint8_t *>(p);
      slots[(size_t)rank].assign(b, b + n);
    }
    barrier();
    std::vector<std::vector<uint8_t>> all;
    {
      std::lock_guard<std::mutex> lk(mu);
      all = slots;
    }
    barrier();
    return all;
  }
};
// the rank thread's hub: mpi::AllReduce has no context argument (MPI_COMM_WORLD in the reference)
thread_local MemHub *t_hub = nullptr;
thread_local int t_rank = 0;

class MemChannel : public cylon::Channel {
 public:
  MemChannel(MemHub *hub, int rank) : hub_(hub), rank_(rank) {}
  void init(int edge, const std::vector<int> &receives, const std::vector<int> &sendIds, cylon::ChannelReceiveCallback *rcv,
            cylon::ChannelSendCallback *send, cylon::Allocator *alloc) override {
    edge_ = edge;
    sources_ = receives;
    rcv_ = rcv;
    send_ = send;
    alloc_ = alloc;
    for (int t : sendIds) out_[t];
  }
  // Sends follow MPIChannel's per-target state machine (net/mpi/mpi_channel.cpp:139-205), one transition per
  // progressSends() call, with transfers that complete at once: data of a target goes out in the order it was accepted,
  // the finish marker only when no data is waiting, and a completion is reported in the call AFTER the one that posted
  // the message. AllToAll::sendFinishComplete relies on that pacing (it marks sends[target] finished, which is a different
  // entry than the one that sent the marker whenever rank != 0).
  int send(std::shared_ptr<cylon::TxRequest> r) override {
    out_[r->target].waiting.push_back(r);
    return 1;
  }
  int sendFin(std::shared_ptr<cylon::TxRequest> r) override {
    auto &t = out_[r->target];
    if (t.fin) return -1;
    t.fin = r;
    return 1;
  }
  void post(int target, const std::shared_ptr<cylon::TxRequest> &r, bool fin) {
    MemMsg m;
    m.fin = fin;
    if (!fin) {
      m.header.assign(r->header, r->header + r->headerLength);
      const uint8_t *b = static_cast<const uint8_t *>(r->buffer);
      if (r->length > 0) m.data.assign(b, b + r->length);
    }
    std::lock_guard<std::mutex> lk(hub_->mu);
    hub_->queues[{edge_, rank_, target}].push_back(std::move(m));
  }
  void progressSends() override {
    for (auto &kv : out_) {
      auto &t = kv.second;
      switch (t.state) {
        case 0:  // idle
          if (!t.waiting.empty()) {
            t.current = t.waiting.front();
            t.waiting.pop_front();
            post(kv.first, t.current, false);
            t.state = 1;
          } else if (t.fin) {
            post(kv.first, t.fin, true);
            t.state = 2;
          }
          break;
        case 1: {  // a data message is out: report it, post what comes next
          auto done = t.current;
          t.current = nullptr;
          if (!t.waiting.empty()) {
            t.current = t.waiting.front();
            t.waiting.pop_front();
            post(kv.first, t.current, false);
            send_->sendComplete(done);
          } else {
            send_->sendComplete(done);
            if (t.fin) {
              post(kv.first, t.fin, true);
              t.state = 2;
            } else {
              t.state = 0;
            }
          }
          break;
        }
        case 2:  // the finish marker is out
          send_->sendFinishComplete(t.fin);
          t.state = 3;
          break;
        default: break;
      }
    }
  }
  void progressReceives() override {
    for (int src : sources_) {
      while (true) {
        MemMsg m;
        {
          std::lock_guard<std::mutex> lk(hub_->mu);
          auto &q = hub_->queues[{edge_, src, rank_}];
          if (q.empty()) break;
          m = std::move(q.front());
          q.pop_front();
        }
        if (m.fin) {
          rcv_->receivedHeader(src, 1, nullptr, 0);
          continue;
        }
        int *h = nullptr;
        if (!m.header.empty()) {
          h = new int[m.header.size()];
          std::memcpy(h, m.header.data(), m.header.size() * sizeof(int));
        }
        rcv_->receivedHeader(src, 0, h, (int)m.header.size());
        std::shared_ptr<cylon::Buffer> buf;
        alloc_->Allocate((int64_t)m.data.size(), &buf);
        if (!m.data.empty()) std::memcpy(buf->GetByteBuffer(), m.data.data(), m.data.size());
        rcv_->receivedData(src, buf, (int)m.data.size());
      }
    }
  }
  void close() override {}

 private:
  MemHub *hub_;
  int rank_, edge_ = 0;
  std::vector<int> sources_;
  cylon::ChannelReceiveCallback *rcv_ = nullptr;
  cylon::ChannelSendCallback *send_ = nullptr;
  cylon::Allocator *alloc_ = nullptr;
  struct Out {
    std::deque<std::shared_ptr<cylon::TxRequest>> waiting;
    std::shared_ptr<cylon::TxRequest> current, fin;
    int state = 0;
  };
  std::map<int, Out> out_;
};

class MemCommunicator : public cylon::net::Communicator {
 public:
  MemCommunicator(MemHub *hub, int r) : hub_(hub) {
    rank = r;
    world_size = hub->world;
  }
  void Init(const std::shared_ptr<cylon::net::CommConfig> &) override {}
  cylon::Channel *CreateChannel() override { return new MemChannel(hub_, rank); }
  int GetRank() override { return rank; }
  int GetWorldSize() override { return world_size; }
  void Finalize() override {}
  void Barrier() override { hub_->barrier(); }
  cylon::net::CommType GetCommType() override { return cylon::net::CommType::MPI; }

 private:
  MemHub *hub_;
};

// element-wise reduction over what every rank deposited (net/mpi/mpi_operations.cpp:24-90 maps the same DataTypes and ops onto MPI)
template <typename T>
void reduce_ranks(const std::vector<std::vector<uint8_t>> &all, void *out, int count, cylon::net::ReduceOp op) {
  T *o = static_cast<T *>(out);
  for (int i = 0; i < count; i++) {
    T acc;
    std::memcpy(&acc, all[0].data() + (size_t)i * sizeof(T), sizeof(T));
    for (size_t r = 1; r < all.size(); r++) {
      T v;
      std::memcpy(&v, all[r].data() + (size_t)i * sizeof(T), sizeof(T));
      switch (op) {
        case cylon::net::SUM: acc = (T)(acc + v); break;
        case cylon::net::MIN: acc = v < acc ? v : acc; break;
        case cylon::net::MAX: acc = v > acc ? v : acc; break;
      }
    }
    o[i] = acc;
  }
}

template <typename ARROW_T>
cylon::Status minmax_typed(const arrow::Datum &input, std::shared_ptr<cylon::compute::Result> &output) {
  using S = typename arrow::TypeTraits<ARROW_T>::ScalarType;
  using A = typename arrow::TypeTraits<ARROW_T>::ArrayType;
  using T = typename ARROW_T::c_type;
  arrow::ArrayVector chunks;
  if (input.is_chunked_array()) chunks = input.chunked_array()->chunks();
  else chunks.push_back(input.make_array());
  // nulls skipped (MinMaxOptions::SKIP in the reference); an empty input gives (max, lowest) like arrow's kernel state
  T mm[2] = {std::numeric_limits<T>::max(), std::numeric_limits<T>::lowest()};
  for (const auto &c : chunks) {
    const auto arr = std::static_pointer_cast<A>(c);
    for (int64_t i = 0; i < arr->length(); i++) {
      if (arr->IsNull(i)) continue;
      const T v = arr->Value(i);
      if (v < mm[0]) mm[0] = v;
      if (v > mm[1]) mm[1] = v;
    }
  }
  if (t_hub && t_hub->world > 1) {
    const auto all = t_hub->exchange(t_rank, mm, sizeof(mm));
    for (const auto &b : all) {
      T v[2];
      std::memcpy(v, b.data(), sizeof(v));
      if (v[0] < mm[0]) mm[0] = v[0];
      if (v[1] > mm[1]) mm[1] = v[1];
    }
  }
  auto st = arrow::StructScalar::Make({std::make_shared<S>(mm[0]), std::make_shared<S>(mm[1])}, {"min", "max"});
  if (!st.ok()) return cylon::Status(cylon::Code::ExecutionError, st.status().message());
  output = std::make_shared<cylon::compute::Result>(arrow::Datum(st.ValueOrDie()));
  return cylon::Status::OK();
}

}  // namespace

// ---- stand-ins for the two MPI-backed functions DistributedSort's range partitioning calls (NOT reference code: the
// reference's compute/aggregates.cpp and net/mpi/mpi_operations.cpp need MPI). Same contract, ranks = threads of this process.
namespace cylon {
namespace mpi {
cylon::Status AllReduce(const void *send_buf, void *rcv_buf, int count, const std::shared_ptr<DataType> &data_type, cylon::net::ReduceOp op) {
  int w = 0;
  switch (data_type->getType()) {
    case Type::UINT8: case Type::INT8: w = 1; break;
    case Type::UINT16: case Type::INT16: w = 2; break;
    case Type::UINT32: case Type::INT32: case Type::FLOAT: w = 4; break;
    case Type::UINT64: case Type::INT64: case Type::DOUBLE: w = 8; break;
    default: return cylon::Status(cylon::Code::NotImplemented, "AllReduce: type not supported by the checker");
  }
  std::vector<std::vector<uint8_t>> all;
  if (t_hub && t_hub->world > 1) {
    all = t_hub->exchange(t_rank, send_buf, (size_t)count * w);
  } else {
    const uint8_t *b = static_cast<const uint8_t *>(send_buf);
    all.emplace_back(b, b + (size_t)count * w);
  }
  switch (data_type->getType()) {
    case Type::UINT8: reduce_ranks<uint8_t>(all, rcv_buf, count, op); break;
    case Type::INT8: reduce_ranks<int8_t>(all, rcv_buf, count, op); break;
    case Type::UINT16: reduce_ranks<uint16_t>(all, rcv_buf, count, op); break;
    case Type::INT16: reduce_ranks<int16_t>(all, rcv_buf, count, op); break;
    case Type::UINT32: reduce_ranks<uint32_t>(all, rcv_buf, count, op); break;
    case Type::INT32: reduce_ranks<int32_t>(all, rcv_buf, count, op); break;
    case Type::UINT64: reduce_ranks<uint64_t>(all, rcv_buf, count, op); break;
    case Type::INT64: reduce_ranks<int64_t>(all, rcv_buf, count, op); break;
    case Type::FLOAT: reduce_ranks<float>(all, rcv_buf, count, op); break;
    default: reduce_ranks<double>(all, rcv_buf, count, op); break;
  }
  return cylon::Status::OK();
}
}  // namespace mpi
namespace compute {
// local min / max (nulls skipped), then min / max over the ranks -- what compute/aggregates.cpp:82-152 does
cylon::Status MinMax(std::shared_ptr<cylon::CylonContext> &, const arrow::Datum &array, const std::shared_ptr<cylon::DataType> &,
                     std::shared_ptr<cylon::compute::Result> &output) {
  const arrow::Datum &local = array;
  switch (array.type()->id()) {
    case arrow::Type::UINT8: return minmax_typed<arrow::UInt8Type>(local, output);
    case arrow::Type::INT8: return minmax_typed<arrow::Int8Type>(local, output);
    case arrow::Type::UINT16: return minmax_typed<arrow::UInt16Type>(local, output);
    case arrow::Type::INT16: return minmax_typed<arrow::Int16Type>(local, output);
    case arrow::Type::UINT32: return minmax_typed<arrow::UInt32Type>(local, output);
    case arrow::Type::INT32: return minmax_typed<arrow::Int32Type>(local, output);
    case arrow::Type::UINT64: return minmax_typed<arrow::UInt64Type>(local, output);
    case arrow::Type::INT64: return minmax_typed<arrow::Int64Type>(local, output);
    case arrow::Type::FLOAT: return minmax_typed<arrow::FloatType>(local, output);
    case arrow::Type::DOUBLE: return minmax_typed<arrow::DoubleType>(local, output);
    default: return cylon::Status(cylon::Code::NotImplemented, "MinMax: type not supported by the checker");
  }
}
}  // namespace compute
}  // namespace cylon

extern "C" {

// One Distributed* operator of the reference on `world` ranks (threads). lcols / rcols: world x nlc / world x nrc column
// descriptors, rank-major. op: 0 DistributedJoin(p1 = left keys, p2 = right keys, join_type, algorithm, prefixes l_/r_),
// 1/2/3 DistributedUnion/Subtract/Intersect, 4 DistributedUnique(p1 = columns), 5 Shuffle(p1 = hash columns),
// 6 DistributedHashGroupBy(p1 = index columns, p2 = (aggregate column, AggregationOpId) pairs flattened),
// 7 DistributedSort(p1 = sort columns, p2 = directions (1 = ascending), join_type = num_bins, algorithm = num_samples; -1 = the
//   rank's row count, i.e. no sampling), 8 MapToSortPartitions(column p1[0], ascending = p2[0], same options) -> one uint32
//   column of target partitions per rank, 9 Sort(p1 = sort columns, p2 = directions), rank-local.
// out: `world` result tables (ref_table_*), in rank order. Returns the first non-zero cylon::Code, or 0.
int ref_dist_op(int32_t op, int32_t world, const ref_col_in *lcols, int32_t nlc, const ref_col_in *rcols, int32_t nrc, const int32_t *p1,
                int32_t n1, const int32_t *p2, int32_t n2, int32_t join_type, int32_t algorithm, ref_table **out) {
  MemHub hub;
  hub.world = world;
  std::vector<int> codes((size_t)world, 0);
  std::vector<std::shared_ptr<arrow::Table>> results((size_t)world);
  auto body = [&](int r) {
    t_hub = &hub;
    t_rank = r;
    auto ctx = std::make_shared<cylon::CylonContext>(world > 1);
    if (world > 1) ctx->setCommunicator(std::make_shared<MemCommunicator>(&hub, r));
    std::shared_ptr<cylon::Table> left, right, res;
    auto lt = make_table(lcols + (size_t)r * nlc, nlc);
    if (!lt || !cylon::Table::FromArrowTable(ctx, lt, left).is_ok()) {
      codes[r] = 42;
      return;
    }
    if (nrc > 0) {
      auto rt = make_table(rcols + (size_t)r * nrc, nrc);
      if (!rt || !cylon::Table::FromArrowTable(ctx, rt, right).is_ok()) {
        codes[r] = 42;
        return;
      }
    }
    cylon::Status st;
    const std::vector<int> v1(p1, p1 + n1);
    switch (op) {
      case 0: {
        cylon::join::config::JoinConfig cfg(static_cast<cylon::join::config::JoinType>(join_type), v1, std::vector<int>(p2, p2 + n2),
                                            static_cast<cylon::join::config::JoinAlgorithm>(algorithm), "l_", "r_");
        st = cylon::DistributedJoin(left, right, cfg, res);
        break;
      }
      case 1: st = cylon::DistributedUnion(left, right, res); break;
      case 2: st = cylon::DistributedSubtract(left, right, res); break;
      case 3: st = cylon::DistributedIntersect(left, right, res); break;
      case 4: st = cylon::DistributedUnique(left, v1, res); break;
      case 5: st = cylon::Shuffle(left, v1, res); break;
      case 6: {
        std::vector<int32_t> agg_cols;
        std::vector<cylon::compute::AggregationOpId> agg_ops;
        for (int i = 0; i + 1 < n2; i += 2) {
          agg_cols.push_back(p2[i]);
          agg_ops.push_back(static_cast<cylon::compute::AggregationOpId>(p2[i + 1]));
        }
        st = cylon::DistributedHashGroupBy(left, std::vector<int32_t>(p1, p1 + n1), agg_cols, agg_ops, res);
        break;
      }
      case 7:
      case 8:
      case 9: {
        std::vector<bool> dirs;
        for (int i = 0; i < n2; i++) dirs.push_back(p2[i] != 0);
        const uint64_t samples = algorithm < 0 ? (uint64_t)left->Rows() : (uint64_t)algorithm;
        if (op == 7) {
          st = cylon::DistributedSort(left, v1, res, dirs, cylon::SortOptions{(uint32_t)join_type, samples});
        } else if (op == 9) {
          st = cylon::Sort(left, std::vector<int32_t>(v1.begin(), v1.end()), res, dirs);
        } else {
          std::vector<uint32_t> target, hist;
          st = cylon::MapToSortPartitions(left, v1[0], (uint32_t)world, target, hist, dirs[0], samples, (uint32_t)join_type);
          if (st.is_ok()) {
            arrow::UInt32Builder b;
            (void)b.AppendValues(target);
            std::shared_ptr<arrow::Array> arr;
            (void)b.Finish(&arr);
            auto t = arrow::Table::Make(arrow::schema({arrow::field("pid", arrow::uint32())}), {arr});
            st = cylon::Table::FromArrowTable(ctx, t, res);
          }
        }
        break;
      }
      default: st = cylon::Status(cylon::Code::Invalid, "unknown op");
    }
    if (!st.is_ok()) {
      fprintf(stderr, "ref_dist_op(rank %d): %s\n", r, st.get_msg().c_str());
      codes[r] = st.get_code() ? st.get_code() : 42;
      return;
    }
    auto comb = res->get_table()->CombineChunks();
    if (!comb.ok()) {
      codes[r] = 42;
      return;
    }
    results[r] = comb.ValueOrDie();
  };
  if (world == 1) {
    body(0);
  } else {
    std::vector<std::thread> threads;
    for (int r = 0; r < world; r++) threads.emplace_back(body, r);
    for (auto &t : threads) t.join();
  }
  for (int r = 0; r < world; r++)
    if (codes[r]) return codes[r];
  for (int r = 0; r < world; r++) out[r] = new ref_table{results[r]};
  return 0;
}

}  // extern "C"
