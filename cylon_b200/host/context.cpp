// context.cpp -- cylon::CylonContext and the NCCL communicator.
//
// Replaces (reference, relative to cpp/src/cylon/): ctx/cylon_context.cpp:24-108 and the transport stack
// net/mpi/mpi_communicator.cpp:40-70, net/mpi/mpi_channel.cpp:30-247, net/ops/all_to_all.cpp:26-178. The reference
// moves every Arrow buffer with a header Isend, a payload Isend and a FIN message per peer, polled with MPI_Test
// from a busy loop. Here one process drives one GPU and a table shuffle is a single ncclGroupStart ...
// ncclSend/ncclRecv x (P-1) x buffers ... ncclGroupEnd enqueued on the context's stream; NVSwitch gives every pair
// of GPUs the full 900 GB/s per direction so the flat all-to-all needs no topology staging.
// libnccl is resolved with dlopen at first use so that the library also loads on hosts without NCCL/GPU
// (the build container); inside a torch process this binds to the libnccl.so.2 torch already loaded.
#include "cylon/ctx/cylon_context.hpp"

#include <dlfcn.h>
#include <nccl.h>

#include <condition_variable>
#include <cstdlib>
#include <iterator>
#include <map>
#include <set>
#include <mutex>

#include "../csrc/common.cuh"
#include "util.hpp"

namespace cylon {

namespace {

struct NcclApi {
  void *handle = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclSend) Send = nullptr;
  decltype(&ncclRecv) Recv = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  std::string error;
};

NcclApi *LoadNccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
      api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
    if (!api.handle) {
      api.error = std::string("cannot load libnccl.so.2: ") + dlerror();
      return;
    }
#define LOAD(field, sym)                                                   \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, sym)); \
  if (!api.field) api.error = std::string("libnccl lacks symbol ") + sym;
    LOAD(GetUniqueId, "ncclGetUniqueId")
    LOAD(CommInitRank, "ncclCommInitRank")
    LOAD(CommDestroy, "ncclCommDestroy")
    LOAD(GroupStart, "ncclGroupStart")
    LOAD(GroupEnd, "ncclGroupEnd")
    LOAD(Send, "ncclSend")
    LOAD(Recv, "ncclRecv")
    LOAD(AllGather, "ncclAllGather")
    LOAD(AllReduce, "ncclAllReduce")
    LOAD(GetErrorString, "ncclGetErrorString")
#undef LOAD
  });
  return &api;
}

#define NCCL_RETURN_IF_FAILED(api, expr)                                                                        \
  do {                                                                                                          \
    ncclResult_t _r = (expr);                                                                                   \
    if (_r != ncclSuccess) return Status(Code::ExecutionError, std::string("NCCL error: ") + (api)->GetErrorString(_r) + " in " #expr); \
  } while (0)

class NcclCommunicator : public net::Communicator {
 public:
  NcclCommunicator(NcclApi *api, ncclComm_t comm, int rank, int world) : api_(api), comm_(comm), rank_(rank), world_(world) {}

  Status StreamBarrier(void *stream) override {
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!barrier_word_) CUDA_RETURN_IF_FAILED(cudaMalloc(&barrier_word_, 64));
    NCCL_RETURN_IF_FAILED(api_, api_->AllReduce(barrier_word_, static_cast<char *>(barrier_word_) + 32, 1, ncclInt32, ncclSum, comm_, s));
    return Status::OK();
  }

  // Receive buffers of the other ranks, mapped through CUDA IPC so that kernels here can store into them over NVLink.
  // Handles are cached: the allocator recycles the same blocks step after step, so after warm-up nothing is opened.
  Status MapPeerBuffers(void *const *mine, int n, void **peers, bool *all_ok, void *stream) override {
    constexpr int kWords = sizeof(cudaIpcMemHandle_t) / 8;  // 64-byte handle = 8 words
    const int per_rank = n * kWords + 2;                    // + "my handles are fine" + allocator generation
    std::vector<int64_t> send((size_t)per_rank), recv((size_t)per_rank * world_);
    bool ok = !ipc_broken_;
    for (int i = 0; i < n && ok; i++) {
      cudaIpcMemHandle_t h;
      if (cudaIpcGetMemHandle(&h, mine[i]) != cudaSuccess) {
        (void)cudaGetLastError();
        ok = false;
        break;
      }
      memcpy(&send[(size_t)i * kWords], &h, sizeof(h));
    }
    send[(size_t)n * kWords] = ok ? 1 : 0;
    send[(size_t)n * kWords + 1] = (int64_t)cylb::pool_generation();
    RETURN_CYLON_STATUS_IF_FAILED(AllGather(send.data(), per_rank, recv.data(), stream));
    if (peer_generation_.size() != (size_t)world_) peer_generation_.assign((size_t)world_, -1);
    // `any_new` is computed from the gathered data alone, so every rank reaches the same value: it decides whether the
    // agreement exchange below runs, and a collective must be entered by all ranks or by none.
    bool any_new = false;
    for (int r = 0; r < world_; r++) {
      // A rank that trimmed its allocator since the last exchange may have given the same addresses to other blocks:
      // every mapping of its memory cached anywhere is closed and opened again from the handles it sends now.
      const int64_t gen = recv[(size_t)r * per_rank + (size_t)n * kWords + 1];
      if (peer_generation_[r] != gen) {
        for (auto it = seen_.begin(); it != seen_.end();) it = it->first == r ? seen_.erase(it) : std::next(it);
        for (auto it = ipc_cache_.begin(); it != ipc_cache_.end();) {
          if (it->first.first == r) {
            cudaIpcCloseMemHandle(it->second);
            it = ipc_cache_.erase(it);
          } else {
            ++it;
          }
        }
        peer_generation_[r] = gen;
      }
      for (int i = 0; i < n; i++) {
        std::pair<int, std::string> key(r, std::string(reinterpret_cast<const char *>(&recv[(size_t)r * per_rank + (size_t)i * kWords]),
                                                       sizeof(cudaIpcMemHandle_t)));
        if (seen_.insert(key).second) any_new = true;
      }
    }
    for (int r = 0; r < world_ && ok; r++) {
      if (r == rank_) {
        for (int i = 0; i < n; i++) peers[(size_t)r * n + i] = mine[i];
        continue;
      }
      for (int i = 0; i < n && ok; i++) {
        std::pair<int, std::string> key(r, std::string(reinterpret_cast<const char *>(&recv[(size_t)r * per_rank + (size_t)i * kWords]),
                                                       sizeof(cudaIpcMemHandle_t)));
        auto it = ipc_cache_.find(key);
        if (it == ipc_cache_.end()) {
          cudaIpcMemHandle_t h;
          memcpy(&h, key.second.data(), sizeof(h));
          void *p = nullptr;
          if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            (void)cudaGetLastError();
            ok = false;
            break;
          }
          it = ipc_cache_.emplace(key, p).first;
        }
        peers[(size_t)r * n + i] = it->second;
      }
    }
    // Agreement: one rank failing to map means nobody uses peer stores. Mapping can only fail for a handle that is new,
    // so the exchange is skipped when every handle of this call was seen before (the allocator recycles the same blocks
    // step after step) and every rank had reported its own handles fine.
    bool all_fine = true;
    for (int r = 0; r < world_; r++) all_fine = all_fine && recv[(size_t)r * per_rank + (size_t)n * kWords] != 0;
    if (any_new || !all_fine || ipc_broken_) {
      int64_t mine_ok = ok ? 1 : 0;
      std::vector<int64_t> oks((size_t)world_);
      RETURN_CYLON_STATUS_IF_FAILED(AllGather(&mine_ok, 1, oks.data(), stream));
      *all_ok = true;
      for (int64_t v : oks) *all_ok = *all_ok && v != 0;
    } else {
      *all_ok = true;
    }
    if (!*all_ok) ipc_broken_ = true;
    return Status::OK();
  }
  ~NcclCommunicator() override { Finalize(); }
  int GetRank() const override { return rank_; }
  int GetWorldSize() const override { return world_; }
  net::CommType GetCommType() const override { return net::CommType::NCCL; }

  void Barrier() override {
    int64_t v = rank_;
    std::vector<int64_t> all((size_t)world_);
    AllGather(&v, 1, all.data(), nullptr);
  }

  void Finalize() override {
    for (auto &kv : ipc_cache_) cudaIpcCloseMemHandle(kv.second);
    ipc_cache_.clear();
    if (barrier_word_) cudaFree(barrier_word_);
    barrier_word_ = nullptr;
    if (comm_) {
      api_->CommDestroy(comm_);
      comm_ = nullptr;
    }
  }

  Status AllGather(const int64_t *send, int n, int64_t *recv, void *stream) override {
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    detail::DevBuf ds, dr;
    RETURN_CYLON_STATUS_IF_FAILED(ds.alloc((size_t)n * 8, stream));
    RETURN_CYLON_STATUS_IF_FAILED(dr.alloc((size_t)n * 8 * world_, stream));
    CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(ds.p, send, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    NCCL_RETURN_IF_FAILED(api_, api_->AllGather(ds.p, dr.p, (size_t)n, ncclInt64, comm_, s));
    CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(recv, dr.p, (size_t)n * 8 * world_, cudaMemcpyDeviceToHost, s));
    CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(s));
    return Status::OK();
  }

  Status AllToAllV(const std::vector<Exchange> &ex, void *stream) override {
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    cylb::ProfScope ps("shuffle_alltoall", s);
    // own slice: plain device copy, enqueued first so it overlaps the NVLink traffic
    for (const auto &e : ex) {
      const int64_t b = e.send_bytes[rank_];
      if (b > 0)
        CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(static_cast<char *>(e.recv) + e.recv_off[rank_], static_cast<const char *>(e.send) + e.send_off[rank_],
                                              (size_t)b, cudaMemcpyDeviceToDevice, s));
    }
    NCCL_RETURN_IF_FAILED(api_, api_->GroupStart());
    for (const auto &e : ex) {
      for (int p = 0; p < world_; p++) {
        if (p == rank_) continue;
        if (e.send_bytes[p] > 0)
          NCCL_RETURN_IF_FAILED(api_, api_->Send(static_cast<const char *>(e.send) + e.send_off[p], (size_t)e.send_bytes[p], ncclChar, p, comm_, s));
        if (e.recv_bytes[p] > 0)
          NCCL_RETURN_IF_FAILED(api_, api_->Recv(static_cast<char *>(e.recv) + e.recv_off[p], (size_t)e.recv_bytes[p], ncclChar, p, comm_, s));
      }
    }
    NCCL_RETURN_IF_FAILED(api_, api_->GroupEnd());
    return Status::OK();
  }

 private:
  NcclApi *api_;
  ncclComm_t comm_;
  int rank_, world_;
  void *barrier_word_ = nullptr;
  bool ipc_broken_ = false;
  std::map<std::pair<int, std::string>, void *> ipc_cache_;  // (peer rank, IPC handle) -> mapping in this process
  std::vector<int64_t> peer_generation_;                      // allocator generation each peer reported last
  std::set<std::pair<int, std::string>> seen_;                // every (rank, handle) exchanged so far, identical on all ranks
};

// ---- ranks as threads of one process (net::ThreadConfig) ---------------------------------------------------------------
struct ThreadGroup {
  explicit ThreadGroup(int world) : world(world), gathered((size_t)world), exchanges((size_t)world), pointers((size_t)world) {}
  const int world;
  std::mutex mu;
  std::condition_variable cv;
  int waiting = 0;
  long long generation = 0;
  std::vector<std::vector<int64_t>> gathered;                                  // AllGather contributions by rank
  std::vector<const std::vector<net::Communicator::Exchange> *> exchanges;     // AllToAllV descriptors by rank
  std::vector<std::vector<void *>> pointers;                                   // MapPeerBuffers contributions by rank
  void Barrier() {
    std::unique_lock<std::mutex> lk(mu);
    const long long gen = generation;
    if (++waiting == world) {
      waiting = 0;
      generation++;
      cv.notify_all();
    } else {
      cv.wait(lk, [&] { return generation != gen; });
    }
  }
};

std::shared_ptr<ThreadGroup> FindThreadGroup(const std::string &name, int world) {
  static std::mutex mu;
  static std::map<std::string, std::weak_ptr<ThreadGroup>> groups;
  std::lock_guard<std::mutex> lk(mu);
  auto g = groups[name].lock();
  if (!g || g->world != world) {
    g = std::make_shared<ThreadGroup>(world);
    groups[name] = g;
  }
  return g;
}

class ThreadCommunicator : public net::Communicator {
 public:
  ThreadCommunicator(std::shared_ptr<ThreadGroup> group, int rank) : g_(std::move(group)), rank_(rank) {}
  int GetRank() const override { return rank_; }
  int GetWorldSize() const override { return g_->world; }
  net::CommType GetCommType() const override { return net::CommType::THREADS; }
  void Barrier() override { g_->Barrier(); }
  void Finalize() override {}

  Status AllGather(const int64_t *send, int n, int64_t *recv, void *) override {
    g_->gathered[(size_t)rank_].assign(send, send + n);
    g_->Barrier();  // every contribution is in place
    for (int r = 0; r < g_->world; r++) memcpy(recv + (size_t)r * n, g_->gathered[(size_t)r].data(), (size_t)n * 8);
    g_->Barrier();  // everybody has read before anybody overwrites
    return Status::OK();
  }

  // Every rank publishes its descriptors once its send buffers are complete, then PULLS its slices out of the peers'
  // buffers with device-to-device copies on its own stream.
  Status AllToAllV(const std::vector<Exchange> &ex, void *stream) override {
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(s));
    g_->exchanges[(size_t)rank_] = &ex;
    g_->Barrier();
    cudaError_t err = cudaSuccess;
    for (int p = 0; p < g_->world && err == cudaSuccess; p++) {
      const auto &theirs = *g_->exchanges[(size_t)p];
      for (size_t e = 0; e < ex.size() && err == cudaSuccess; e++) {
        const int64_t bytes = ex[e].recv_bytes[(size_t)p];
        if (bytes <= 0) continue;
        err = cudaMemcpyAsync(static_cast<char *>(ex[e].recv) + ex[e].recv_off[(size_t)p],
                              static_cast<const char *>(theirs[e].send) + theirs[e].send_off[(size_t)rank_], (size_t)bytes, cudaMemcpyDefault, s);
      }
    }
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);
    g_->Barrier();  // all pulls are done: send buffers may be released
    if (err != cudaSuccess) return Status(Code::ExecutionError, std::string("thread all-to-all: ") + cudaGetErrorString(err));
    return Status::OK();
  }

  Status StreamBarrier(void *stream) override {
    CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    g_->Barrier();
    return Status::OK();
  }

  // same process: a peer's allocation is addressable as it is (same device, or peer access / unified addressing across devices)
  Status MapPeerBuffers(void *const *mine, int n, void **peers, bool *all_ok, void *) override {
    g_->pointers[(size_t)rank_].assign(mine, mine + n);
    g_->Barrier();
    for (int r = 0; r < g_->world; r++)
      for (int i = 0; i < n; i++) peers[(size_t)r * n + i] = g_->pointers[(size_t)r][(size_t)i];
    g_->Barrier();
    *all_ok = true;
    return Status::OK();
  }

 private:
  std::shared_ptr<ThreadGroup> g_;
  int rank_;
};

thread_local std::string g_ctx_error;

}  // namespace

Status net::NcclConfig::CreateUniqueId(NcclUniqueId *out) {
  NcclApi *api = LoadNccl();
  if (!api->error.empty()) return Status(Code::ExecutionError, api->error);
  static_assert(sizeof(ncclUniqueId) == sizeof(NcclUniqueId), "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  NCCL_RETURN_IF_FAILED(api, api->GetUniqueId(&id));
  memcpy(out->bytes, &id, sizeof(id));
  return Status::OK();
}

const std::string &CylonContext::LastError() { return g_ctx_error; }

std::shared_ptr<CylonContext> CylonContext::Init() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  return Init(dev);
}

std::shared_ptr<CylonContext> CylonContext::Init(int device) {
  if (cyl_init(device) != CYL_OK) {
    g_ctx_error = cyl_last_error();
    return nullptr;
  }
  auto ctx = std::make_shared<CylonContext>(false);
  ctx->device = device;
  return ctx;
}

std::shared_ptr<CylonContext> CylonContext::InitDistributed(const std::shared_ptr<net::CommConfig> &config) {
  if (config && config->Type() == net::CommType::THREADS) {
    auto tc = std::static_pointer_cast<net::ThreadConfig>(config);
    if (tc->world_size < 1 || tc->rank < 0 || tc->rank >= tc->world_size) {
      g_ctx_error = "InitDistributed: ThreadConfig rank / world size out of range";
      return nullptr;
    }
    if (cyl_init(tc->device) != CYL_OK) {
      g_ctx_error = cyl_last_error();
      return nullptr;
    }
    auto ctx = std::make_shared<CylonContext>(true);
    ctx->device = tc->device;
    ctx->communicator = std::make_shared<ThreadCommunicator>(FindThreadGroup(tc->group, tc->world_size), tc->rank);
    return ctx;
  }
  if (!config || config->Type() != net::CommType::NCCL) {
    g_ctx_error = "InitDistributed: pass a cylon::net::NcclConfig (one process per GPU) or a cylon::net::ThreadConfig (ranks as threads)";
    return nullptr;
  }
  auto nc = std::static_pointer_cast<net::NcclConfig>(config);
  if (cyl_init(nc->device) != CYL_OK) {
    g_ctx_error = cyl_last_error();
    return nullptr;
  }
  NcclApi *api = LoadNccl();
  if (!api->error.empty()) {
    g_ctx_error = api->error;
    return nullptr;
  }
  // The shuffle is pure point-to-point traffic. NCCL's default gives send/recv only a few channels per peer, which
  // leaves NVLink mostly idle when there are few peers (measured at 2 GPUs: 316 GB/s per direction); ask for more
  // unless the user already chose. Must happen before the communicator is created.
  setenv("NCCL_MIN_P2P_NCHANNELS", "32", 0);
  setenv("NCCL_MAX_P2P_NCHANNELS", "64", 0);
  ncclUniqueId id;
  memcpy(&id, nc->id.bytes, sizeof(id));
  ncclComm_t comm = nullptr;
  ncclResult_t r = api->CommInitRank(&comm, nc->world_size, id, nc->rank);
  if (r != ncclSuccess) {
    g_ctx_error = std::string("ncclCommInitRank failed: ") + api->GetErrorString(r);
    return nullptr;
  }
  auto ctx = std::make_shared<CylonContext>(true);
  ctx->device = nc->device;
  ctx->communicator = std::make_shared<NcclCommunicator>(api, comm, nc->rank, nc->world_size);
  return ctx;
}

void CylonContext::Finalize() {
  if (communicator) communicator->Finalize();
  if (aux_stream) {
    cudaStreamDestroy(static_cast<cudaStream_t>(aux_stream));
    aux_stream = nullptr;
  }
}

void *CylonContext::GetAuxStream() {
  if (!aux_stream) {
    cudaStream_t s = nullptr;
    if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) {
      (void)cudaGetLastError();
      return nullptr;
    }
    aux_stream = s;
  }
  return aux_stream;
}

std::vector<int> CylonContext::GetNeighbours(bool include_self) {
  std::vector<int> n;
  const int w = GetWorldSize(), r = GetRank();
  for (int i = 0; i < w; i++)
    if (include_self || i != r) n.push_back(i);
  return n;
}

}  // namespace cylon
