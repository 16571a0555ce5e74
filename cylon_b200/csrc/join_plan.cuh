// join_plan.cuh -- host object shared by the hash and sort join front ends (count pass -> materialize pass).
#pragma once
#include "common.cuh"

struct cyl_join_plan {
  int algorithm = CYL_JOIN_HASH;
  int join_type = CYL_JOIN_INNER;
  bool build_right = false;  // hash join: build side is the right table (probe = left)
  bool exact = false;        // key tag == key bits, no verification needed
  int64_t nl = 0, nr = 0;    // rows of the left / right table
  int64_t nb = 0, np = 0;    // build / probe rows (hash); right / left rows (sort: "probe" = left)
  int64_t m_probe = 0;       // rows produced by the probe side
  int64_t m_extra = 0;       // FULL_OUTER / RIGHT: rows appended for unmatched rows of the other side
  cylb::DCols bk, pk;
  cudaStream_t stream = nullptr;
  // hash join state
  void *table = nullptr;      // Slot[cap + 1]
  int64_t cap = 0;
  uint32_t *next = nullptr;   // [nb] chain links
  uint32_t *pslot = nullptr;  // [np] slot of each probe row (NIL = none)
  int64_t *out_off = nullptr; // [np + 1]
  uint8_t *bsurv = nullptr;   // [nb] 1 = build row never matched (FULL_OUTER)
  int64_t *bsurv_pos = nullptr;
  // sort join state
  uint32_t *lperm = nullptr, *rperm = nullptr;  // argsorted row ids
  uint32_t *lo = nullptr, *hi = nullptr;        // [nl] match range [lo, hi) in rperm for sorted-left position
  uint8_t *rsurv = nullptr;                     // [nr] (by sorted position) 1 = right row never matched
  int64_t *rsurv_pos = nullptr;
};

namespace cylb {
void plan_release(cyl_join_plan *p);
int hash_join_materialize(cyl_join_plan *p, int64_t *l_idx, int64_t *r_idx, cudaStream_t s);
int sort_join_materialize(cyl_join_plan *p, int64_t *l_idx, int64_t *r_idx, cudaStream_t s);
template <typename T>
int plan_alloc(T **ptr, size_t count, cudaStream_t s) {
  void *q = nullptr;
  cudaError_t e = pool_malloc(&q, (count ? count : 1) * sizeof(T), s);
  if (e != cudaSuccess) return cuda_fail(e, "pool_malloc(join plan)", __FILE__, __LINE__);
  *ptr = static_cast<T *>(q);
  return CYL_OK;
}
}  // namespace cylb
