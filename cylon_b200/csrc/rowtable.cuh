// rowtable.cuh -- open-addressing "row set" in HBM shared by group-by, unique and the set operations.
//
// The reference uses ska::bytell_hash_map/set keyed by row index with a (row hash, row equality) functor pair
// and relies on "first inserted row wins" (groupby/hash_groupby.cpp:92-126, table.cpp:531-721,923-982). Here a
// slot is { u64 tag ; u64 rep }: rows with equal key tags land in the same slot and `rep` ends up as the MINIMUM
// combined row index (atomicMin), i.e. the first occurrence in input order -- left-table rows order before
// right-table rows because the right table's indices carry bit 62. Exact mode: tag == key bits. Hashed mode:
// tag == 64-bit hash of the key tuple; every row is then verified against its slot's representative with the
// reference's value equality and a mismatch (a true 64-bit collision) raises a flag on which the host repeats the
// operation with another seed, so results never depend on the hash.
// The table starts small (L2-resident when the number of distinct keys is small) and the host grows it x4 when
// an insert pass reports that more than half the slots were claimed.
#pragma once
#include "common.cuh"

namespace cylb {

struct __align__(16) RSlot {
  unsigned long long tag;
  unsigned long long rep;
};

constexpr unsigned long long kRtEmpty = 0xFFFFFFFFFFFFFFFFull;
constexpr unsigned long long kRightBit = 1ull << 62;
constexpr uint32_t kNoSlot = 0xFFFFFFFFu;

struct RowTable {
  RSlot *slots;                  // cap + 1 (the last slot serves the tag equal to kRtEmpty in exact mode)
  unsigned long long mask;       // cap - 1
  unsigned long long limit;      // claimed-slot budget (cap / 2)
  unsigned long long *nclaimed;  // device counter
  int *flags;                    // [0] = overflow, [1] = tag collision between unequal rows
};

#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long rt_sanitize(unsigned long long tag, bool exact) {
  return (!exact && tag == kRtEmpty) ? kRtEmpty - 1 : tag;
}

// Finds or claims the slot of `tag`. Returns the slot index, or kNoSlot when the table is full / overflowing.
// *claimed is set when this call claimed a fresh slot.
__device__ __forceinline__ uint32_t rt_find_or_claim(const RowTable &t, unsigned long long tag, bool *claimed) {
  *claimed = false;
  if (tag == kRtEmpty) return (uint32_t)(t.mask + 1);
  unsigned long long s = mix64(tag) & t.mask;
  for (int probes = 0; probes < 8192; probes++) {
    unsigned long long cur = t.slots[s].tag;
    if (cur == tag) return (uint32_t)s;
    if (cur == kRtEmpty) {
      cur = atomicCAS(&t.slots[s].tag, kRtEmpty, tag);
      if (cur == kRtEmpty) {
        *claimed = true;
        return (uint32_t)s;
      }
      if (cur == tag) return (uint32_t)s;
    }
    s = (s + 1) & t.mask;
  }
  return kNoSlot;
}

__device__ __forceinline__ uint32_t rt_find(const RowTable &t, unsigned long long tag) {
  if (tag == kRtEmpty) return (uint32_t)(t.mask + 1);
  unsigned long long s = mix64(tag) & t.mask;
  for (int probes = 0; probes < 8192; probes++) {
    const unsigned long long cur = t.slots[s].tag;
    if (cur == tag) return (uint32_t)s;
    if (cur == kRtEmpty) return kNoSlot;
    s = (s + 1) & t.mask;
  }
  return kNoSlot;
}
#endif

// Host handle: owns the device table, grows it and re-runs the caller's insert pass until it fits.
struct RowTableHost {
  RowTable t{};
  int64_t cap = 0;
  cudaStream_t s = nullptr;
  int alloc(int64_t capacity, cudaStream_t stream);  // allocates + clears
  void release();
  int read_flags(int *overflow, int *collision);      // synchronises
  ~RowTableHost() { release(); }
};

// insert rows [0, n) of `keys` with combined index (row | index_bits); slot_out (optional) gets each row's slot
int rowtable_insert(RowTableHost &rt, const DCols &keys, int64_t n, bool exact, unsigned long long seed, unsigned long long index_bits,
                    uint32_t *slot_out);
int64_t rowtable_initial_capacity(int64_t expected_rows);

}  // namespace cylb
