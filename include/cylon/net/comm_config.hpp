// cylon::net::CommType / CommConfig / Communicator -- the transport seam of the reference
// (cpp/src/cylon/net/comm_type.hpp:20-22, comm_config.hpp:22-36, communicator.hpp:24-37, channel.hpp:51-90),
// with one added backend: NCCL over NVLink (one process per GPU). The reference's only backend is MPI
// point-to-point with a header/payload/FIN handshake per buffer (net/mpi/mpi_channel.cpp); here a table shuffle is
// one count exchange plus ONE grouped ncclSend/ncclRecv over all column buffers (cylon_b200/host/shuffle.cpp).
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../status.hpp"

namespace cylon {
namespace net {

enum CommType { LOCAL = 0, MPI, TCP, UCX, NCCL, THREADS };

class CommConfig {
 public:
  virtual ~CommConfig() = default;
  virtual CommType Type() = 0;
};

// 128-byte NCCL bootstrap id: rank 0 creates it (NcclConfig::CreateUniqueId) and hands it to the other ranks by any
// out-of-band channel (torch.distributed broadcast, a file, MPI_Bcast ...).
struct NcclUniqueId {
  char bytes[128];
};

class NcclConfig : public CommConfig {
 public:
  NcclConfig(int rank, int world_size, int device, const NcclUniqueId &id) : rank(rank), world_size(world_size), device(device), id(id) {}
  static std::shared_ptr<NcclConfig> Make(int rank, int world_size, int device, const NcclUniqueId &id) {
    return std::make_shared<NcclConfig>(rank, world_size, device, id);
  }
  static Status CreateUniqueId(NcclUniqueId *out);
  CommType Type() override { return CommType::NCCL; }
  int rank, world_size, device;
  NcclUniqueId id;
};

// Ranks as host threads of ONE process (each with its own CylonContext and stream, on the same or on different GPUs):
// every thread passes the same group name and world size and its own rank. Counts travel through host memory, payloads as
// device-to-device copies, and the peer-store shuffle writes straight into the other threads' buffers. It exists so that
// everything above the transport -- partition plans, count exchange, offsets, validity and variable-width packing, the
// Distributed* operators -- can be exercised at world > 1 on a box with a single GPU (tests/test_gpu_thread_ranks.py);
// NCCL itself is covered by tests/test_gpu_distributed.py on multi-GPU boxes.
class ThreadConfig : public CommConfig {
 public:
  ThreadConfig(std::string group, int rank, int world_size, int device) : group(std::move(group)), rank(rank), world_size(world_size), device(device) {}
  static std::shared_ptr<ThreadConfig> Make(const std::string &group, int rank, int world_size, int device) {
    return std::make_shared<ThreadConfig>(group, rank, world_size, device);
  }
  CommType Type() override { return CommType::THREADS; }
  std::string group;
  int rank, world_size, device;
};

// What the shuffle needs from a transport. All buffers are device memory; counts and offsets are in bytes.
class Communicator {
 public:
  virtual ~Communicator() = default;
  virtual int GetRank() const = 0;
  virtual int GetWorldSize() const = 0;
  virtual CommType GetCommType() const = 0;
  virtual void Barrier() = 0;
  virtual void Finalize() = 0;
  // every rank contributes n int64 values (host memory); recv gets world*n values ordered by rank
  virtual Status AllGather(const int64_t *send, int n, int64_t *recv, void *stream) = 0;
  struct Exchange {  // one buffer class of one column
    const void *send = nullptr;
    void *recv = nullptr;
    std::vector<int64_t> send_off, send_bytes, recv_off, recv_bytes;  // per peer
  };
  // one grouped all-to-all over every exchange; the own-rank slice is a device-to-device copy
  virtual Status AllToAllV(const std::vector<Exchange> &exchanges, void *stream) = 0;
  // cross-GPU ordering point: work enqueued on `stream` after it, on any rank, starts only after every rank's stream
  // has reached it (and everything those streams did before is complete and visible)
  virtual Status StreamBarrier(void *stream) = 0;
  // Peer memory: mine[0..n) are device allocations of this rank (base pointers from the library's allocator). On
  // success peers[r * n + i] is a pointer THIS GPU can store through to reach rank r's i-th buffer (r == own rank: the
  // local pointer) and *all_ok tells whether every rank succeeded (ranks must agree before relying on it).
  virtual Status MapPeerBuffers(void *const *mine, int n, void **peers, bool *all_ok, void *stream) = 0;
};

}  // namespace net
}  // namespace cylon
