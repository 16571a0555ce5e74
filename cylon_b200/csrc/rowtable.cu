// rowtable.cu -- host side + insert kernel of the shared row set (see rowtable.cuh).
#include "rowtable.cuh"

namespace cylb {

__global__ void rt_clear_kernel(RSlot *__restrict__ s, int64_t n) {
  const uint4 v = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    reinterpret_cast<uint4 *>(s)[i] = v;
}

struct RtInsertArgs {
  DCols keys;
  RowTable t;
  int64_t n;
  int32_t exact;
  unsigned long long seed;
  unsigned long long index_bits;
  uint32_t *slot_out;
};

__global__ void __launch_bounds__(256) rt_insert_kernel(const __grid_constant__ RtInsertArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t npad = (a.n + 31) & ~31ll;  // keep warps converged for the ballot
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npad; i += stride) {
    bool claimed = false;
    if (i < a.n && *reinterpret_cast<volatile int *>(&a.t.flags[0]) == 0) {
      const unsigned long long tag = rt_sanitize(row_tag(a.keys, i, a.exact != 0, a.seed), a.exact != 0);
      const uint32_t s = rt_find_or_claim(a.t, tag, &claimed);
      if (s == kNoSlot) {
        a.t.flags[0] = 1;
      } else {
        atomicMin(&a.t.slots[s].rep, (unsigned long long)i | a.index_bits);
      }
      if (a.slot_out) a.slot_out[i] = s;
    }
    const unsigned newly = __ballot_sync(0xffffffffu, claimed);
    if (lane_id() == 0 && newly) {
      const unsigned long long before = atomicAdd(a.t.nclaimed, (unsigned long long)__popc(newly));
      if (before + __popc(newly) > a.t.limit) a.t.flags[0] = 1;
    }
  }
}

static inline int grid1d(int64_t n) {
  int64_t b = div_up(n, 256);
  const int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

int RowTableHost::alloc(int64_t capacity, cudaStream_t stream) {
  release();
  s = stream;
  cap = capacity;
  void *p = nullptr;
  CYL_CUDA(pool_malloc(&p, (size_t)(cap + 1) * sizeof(RSlot) + 64, s));
  t.slots = static_cast<RSlot *>(p);
  void *q = nullptr;
  CYL_CUDA(pool_malloc(&q, 64, s));
  t.nclaimed = static_cast<unsigned long long *>(q);
  t.flags = reinterpret_cast<int *>(static_cast<char *>(q) + 16);
  t.mask = (unsigned long long)cap - 1;
  t.limit = (unsigned long long)cap / 2;
  CYL_CUDA(cudaMemsetAsync(q, 0, 64, s));
  rt_clear_kernel<<<grid1d(cap + 1), 256, 0, s>>>(t.slots, cap + 1);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

void RowTableHost::release() {
  if (t.slots) pool_free(t.slots, s);
  if (t.nclaimed) pool_free(t.nclaimed, s);
  t.slots = nullptr;
  t.nclaimed = nullptr;
}

int RowTableHost::read_flags(int *overflow, int *collision) {
  int h[2] = {0, 0};
  CYL_CUDA(cudaMemcpyAsync(h, t.flags, sizeof(h), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *overflow = h[0];
  *collision = h[1];
  return CYL_OK;
}

int rowtable_insert(RowTableHost &rt, const DCols &keys, int64_t n, bool exact, unsigned long long seed, unsigned long long index_bits,
                    uint32_t *slot_out) {
  if (n == 0) return CYL_OK;
  RtInsertArgs a;
  memset(&a, 0, sizeof(a));
  a.keys = keys;
  a.t = rt.t;
  a.n = n;
  a.exact = exact ? 1 : 0;
  a.seed = seed;
  a.index_bits = index_bits;
  a.slot_out = slot_out;
  rt_insert_kernel<<<grid1d(n), 256, 0, rt.s>>>(a);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int64_t rowtable_initial_capacity(int64_t expected_rows) {
  // 2^22 slots = 64 MB: stays in the 126 MB L2 when the number of distinct keys is small
  int64_t cap = next_pow2(expected_rows * 2);
  if (cap < 1024) cap = 1024;
  if (cap > (1ll << 22)) cap = 1ll << 22;
  return cap;
}

}  // namespace cylb
