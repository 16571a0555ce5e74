// cylon::CylonContext -- same entry points as the reference (cpp/src/cylon/ctx/cylon_context.hpp:49-145):
// Init(), InitDistributed(config), GetRank(), GetWorldSize(), GetNeighbours(), GetNextSequence(), Barrier(),
// Finalize(), AddConfig/GetConfig. In addition it carries what a GPU rank needs: the CUDA device ordinal and the
// stream every operator of this context is ordered on.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../net/comm_config.hpp"

namespace cylon {

class CylonContext {
 public:
  explicit CylonContext(bool distributed) : is_distributed(distributed) {}
  // local, single-GPU context on the current (or given) device
  static std::shared_ptr<CylonContext> Init();
  static std::shared_ptr<CylonContext> Init(int device);
  // distributed context; config->Type() must be CommType::NCCL (the reference accepts only MPI:
  // ctx/cylon_context.cpp:32-43). Returns nullptr and leaves a message in LastError() on failure.
  static std::shared_ptr<CylonContext> InitDistributed(const std::shared_ptr<net::CommConfig> &config);
  static const std::string &LastError();

  void Finalize();
  void AddConfig(const std::string &key, const std::string &value) { config[key] = value; }
  std::string GetConfig(const std::string &key, const std::string &def = "") {
    auto it = config.find(key);
    return it == config.end() ? def : it->second;
  }
  std::shared_ptr<net::Communicator> GetCommunicator() const { return communicator; }
  void setCommunicator(const std::shared_ptr<net::Communicator> &c) { communicator = c; }
  void setDistributed(bool distributed) { is_distributed = distributed; }
  bool IsDistributed() const { return is_distributed; }
  int GetRank() { return is_distributed && communicator ? communicator->GetRank() : 0; }
  int GetWorldSize() { return is_distributed && communicator ? communicator->GetWorldSize() : 1; }
  std::vector<int> GetNeighbours(bool include_self);
  int32_t GetNextSequence() { return sequence_no++; }
  net::CommType GetCommType() { return is_distributed && communicator ? communicator->GetCommType() : net::CommType::LOCAL; }
  void Barrier() {
    if (IsDistributed() && communicator) communicator->Barrier();
  }
  // GPU additions
  int GetDevice() const { return device; }
  void *GetStream() const { return stream; }   // cudaStream_t
  void SetStream(void *s) { stream = s; }      // caller-owned stream (e.g. torch's current stream)
  void *GetAuxStream();                        // second (non-blocking) stream of this context, created on first use; nullptr on failure

 private:
  std::unordered_map<std::string, std::string> config{};
  bool is_distributed;
  std::shared_ptr<net::Communicator> communicator{};
  int32_t sequence_no = 0;
  int device = 0;
  void *stream = nullptr;
  void *aux_stream = nullptr;
};

}  // namespace cylon
