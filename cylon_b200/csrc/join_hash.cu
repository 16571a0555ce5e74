// join_hash.cu -- local hash join: open-addressing build/probe in HBM, count pass then materialize pass.
//
// Replaces (reference, relative to cpp/src/cylon/): join/hash_join.cpp:21-90 (probe functions), :92-186
// (multi_index_hash_join), :188-307 (ArrayIndexHashJoin). The reference keeps a std::unordered_multimap keyed by
// row index with a precomputed row hash; here the table is an array of 16-byte slots
//     { u64 tag ; u32 head ; u32 count }
// with linear probing on the tag. All build rows with the same tag hang off `head` through a next[] array (the
// multimap's equal_range), so duplicate-heavy build sides cost nothing extra to build and a probe touches
// exactly one slot sector plus the chain it has to emit anyway. When the key is a single fixed-width column the
// tag IS the key (exact mode: slot.count is the match count, no chain walk in the count pass); otherwise the tag
// is a 64-bit hash of the key tuple and every chain entry is verified with the reference's value equality.
// The library's own hash (mix64) places the slots; the reference-exact hash is only needed for partitioning.
#include "join_plan.cuh"
#include "scan.cuh"

namespace cylb {

struct __align__(16) Slot {
  unsigned long long tag;
  uint32_t head;
  uint32_t count;
};

constexpr unsigned long long kEmptyTag = 0xFFFFFFFFFFFFFFFFull;
constexpr uint32_t kNil = 0xFFFFFFFFu;

__global__ void init_slots_kernel(Slot *__restrict__ t, int64_t nslots) {
  const uint4 v = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, kNil, 0u);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nslots; i += (int64_t)gridDim.x * blockDim.x)
    reinterpret_cast<uint4 *>(t)[i] = v;
}

struct HashJoinArgs {
  DCols bk, pk;
  Slot *table;
  int64_t cap_mask;  // cap - 1; slot index `cap` is reserved for the tag that equals kEmptyTag
  uint32_t *next;
  int64_t nb, np;
  int32_t exact;
  int32_t nan_type;     // exact mode with a FLOAT/DOUBLE key: its type id (a NaN key matches nothing), else 0
  int32_t keep_probe;   // LEFT/RIGHT/FULL_OUTER: unmatched probe rows produce one row
  int32_t mark_build;   // FULL_OUTER: clear bsurv[j] for matched build rows
  int32_t build_right;  // orientation of the output pairs
  uint32_t *pslot;
  uint32_t *pcount;
  const int64_t *out_off;
  uint8_t *bsurv;
  const int64_t *bsurv_pos;
  int64_t extra_base;
  int64_t *l_idx, *r_idx;
};

__device__ __forceinline__ unsigned long long sanitize_tag(unsigned long long tag, bool exact) {
  // in hashed mode the (astronomically unlikely) all-ones tag is folded; in exact mode it gets its own slot
  return (!exact && tag == kEmptyTag) ? kEmptyTag - 1 : tag;
}

__global__ void __launch_bounds__(256) hash_build_kernel(const __grid_constant__ HashJoinArgs a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.nb; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long tag = sanitize_tag(row_tag(a.bk, i, a.exact, 0), a.exact);
    if (a.nan_type && key_bits_is_nan(a.nan_type, tag)) {  // never matched: stays a FULL_OUTER survivor, never linked
      a.next[i] = kNil;
      continue;
    }
    int64_t s;
    if (tag == kEmptyTag) {
      s = a.cap_mask + 1;
    } else {
      s = (int64_t)(mix64(tag) & (unsigned long long)a.cap_mask);
      while (true) {
        const unsigned long long old = atomicCAS(&a.table[s].tag, kEmptyTag, tag);
        if (old == kEmptyTag || old == tag) break;
        s = (s + 1) & a.cap_mask;
      }
    }
    a.next[i] = atomicExch(&a.table[s].head, (uint32_t)i);
    atomicAdd(&a.table[s].count, 1u);
  }
}

__device__ __forceinline__ int64_t find_slot(const HashJoinArgs &a, unsigned long long tag) {
  if (tag == kEmptyTag) return a.cap_mask + 1;
  int64_t s = (int64_t)(mix64(tag) & (unsigned long long)a.cap_mask);
  while (true) {
    const unsigned long long t = a.table[s].tag;
    if (t == tag) return s;
    if (t == kEmptyTag) return -1;
    s = (s + 1) & a.cap_mask;
  }
}

__global__ void __launch_bounds__(256) hash_probe_count_kernel(const __grid_constant__ HashJoinArgs a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.np; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long tag = sanitize_tag(row_tag(a.pk, i, a.exact, 0), a.exact);
    const int64_t s = (a.nan_type && key_bits_is_nan(a.nan_type, tag)) ? -1 : find_slot(a, tag);
    uint32_t c = 0;
    if (s >= 0) {
      const uint4 raw = *reinterpret_cast<const uint4 *>(&a.table[s]);
      const uint32_t head = raw.z;
      if (a.exact && !a.mark_build) {
        c = raw.w;
      } else {
        for (uint32_t j = head; j != kNil; j = a.next[j]) {
          if (a.exact || rows_equal(a.bk, j, a.pk, i)) {
            c++;
            if (a.mark_build) a.bsurv[j] = 0;
          }
        }
      }
    }
    a.pslot[i] = c ? (uint32_t)s : kNil;
    a.pcount[i] = c ? c : (a.keep_probe ? 1u : 0u);
  }
}

__global__ void __launch_bounds__(256) hash_probe_write_kernel(const __grid_constant__ HashJoinArgs a) {
  int64_t *build_idx = a.build_right ? a.r_idx : a.l_idx;
  int64_t *probe_idx = a.build_right ? a.l_idx : a.r_idx;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.np; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t o = a.out_off[i];
    const uint32_t s = a.pslot[i];
    if (s == kNil) {
      if (a.keep_probe) {
        build_idx[o] = -1;
        probe_idx[o] = i;
      }
      continue;
    }
    for (uint32_t j = a.table[s].head; j != kNil; j = a.next[j]) {
      if (a.exact || rows_equal(a.bk, j, a.pk, i)) {
        build_idx[o] = j;
        probe_idx[o] = i;
        o++;
      }
    }
  }
}

__global__ void __launch_bounds__(256) hash_build_survivors_kernel(const __grid_constant__ HashJoinArgs a) {
  int64_t *build_idx = a.build_right ? a.r_idx : a.l_idx;
  int64_t *probe_idx = a.build_right ? a.l_idx : a.r_idx;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < a.nb; j += (int64_t)gridDim.x * blockDim.x) {
    if (a.bsurv[j]) {
      const int64_t o = a.extra_base + a.bsurv_pos[j];
      build_idx[o] = j;
      probe_idx[o] = -1;
    }
  }
}

static inline int grid1d(int64_t n) {
  int64_t b = div_up(n, 256);
  const int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

void plan_release(cyl_join_plan *p) {
  if (!p) return;
  void *ptrs[] = {p->table, p->next, p->pslot, p->out_off, p->bsurv, p->bsurv_pos, p->lperm, p->rperm, p->lo, p->hi, p->rsurv, p->rsurv_pos};
  for (void *q : ptrs)
    if (q) pool_free(q, p->stream);
  delete p;
}

static void fill_args(const cyl_join_plan *p, HashJoinArgs *a) {
  memset(a, 0, sizeof(*a));
  a->bk = p->bk;
  a->pk = p->pk;
  a->table = static_cast<Slot *>(p->table);
  a->cap_mask = p->cap - 1;
  a->next = p->next;
  a->nb = p->nb;
  a->np = p->np;
  a->exact = p->exact;
  a->nan_type = (p->exact && (p->bk.c[0].type == CYL_FLOAT || p->bk.c[0].type == CYL_DOUBLE)) ? p->bk.c[0].type : 0;
  a->keep_probe = p->join_type != CYL_JOIN_INNER;
  a->mark_build = p->join_type == CYL_JOIN_FULL_OUTER;
  a->build_right = p->build_right;
  a->pslot = p->pslot;
  a->out_off = p->out_off;
  a->bsurv = p->bsurv;
  a->bsurv_pos = p->bsurv_pos;
}

int hash_join_materialize(cyl_join_plan *p, int64_t *l_idx, int64_t *r_idx, cudaStream_t s) {
  HashJoinArgs a;
  fill_args(p, &a);
  a.l_idx = l_idx;
  a.r_idx = r_idx;
  a.extra_base = p->m_probe;
  if (p->np > 0) {
    ProfScope ps("hash_join_probe_write", s);
    hash_probe_write_kernel<<<grid1d(p->np), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  if (p->m_extra > 0) {
    hash_build_survivors_kernel<<<grid1d(p->nb), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  return CYL_OK;
}

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_hash_join_count(const cyl_column_t *lkeys, const cyl_column_t *rkeys, int32_t nkeys, int64_t nl, int64_t nr, int32_t join_type,
                        cyl_join_plan_t **plan, int64_t *num_rows, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!plan || !num_rows) return fail(CYL_Invalid, "cyl_hash_join_count: null output");
  *plan = nullptr;
  if (nkeys < 1) return fail(CYL_Invalid, "cyl_hash_join_count: needs at least one key column");
  if (join_type < CYL_JOIN_INNER || join_type > CYL_JOIN_FULL_OUTER) return fail(CYL_Invalid, "cyl_hash_join_count: join type %d", join_type);
  for (int c = 0; c < nkeys; c++)
    if (lkeys[c].type != rkeys[c].type)  // hash_join.cpp:195-197
      return fail(CYL_Invalid, "left and right index array types are not equal (key %d: %d vs %d)", c, lkeys[c].type, rkeys[c].type);
  DCols lk, rk;
  CYL_TRY(make_dcols(lkeys, nkeys, nl, &lk, "cyl_hash_join_count(left)"));
  CYL_TRY(make_dcols(rkeys, nkeys, nr, &rk, "cyl_hash_join_count(right)"));

  cyl_join_plan *p = new cyl_join_plan();
  p->algorithm = CYL_JOIN_HASH;
  p->join_type = join_type;
  p->stream = s;
  p->nl = nl;
  p->nr = nr;
  switch (join_type) {  // hash_join.cpp:207-232
    case CYL_JOIN_LEFT: p->build_right = true; break;
    case CYL_JOIN_RIGHT: p->build_right = false; break;
    default: p->build_right = !(nl <= nr); break;
  }
  p->bk = p->build_right ? rk : lk;
  p->pk = p->build_right ? lk : rk;
  p->nb = p->build_right ? nr : nl;
  p->np = p->build_right ? nl : nr;
  p->exact = is_exact_key(lk);
  if (p->nb > (1ll << 30)) {  // slot indices are 32-bit and the table has 2 * nb slots (+ 1 reserved)
    const long long nb = (long long)p->nb;
    delete p;
    return fail(CYL_CapacityError, "cyl_hash_join_count: %lld build rows (max 2^30 per GPU)", nb);
  }
  p->cap = next_pow2(p->nb * 2 < 1024 ? 1024 : p->nb * 2);

#define PLAN_TRY(expr)       \
  do {                       \
    int _rc = (expr);        \
    if (_rc != CYL_OK) {     \
      plan_release(p);       \
      return _rc;            \
    }                        \
  } while (0)
#define PLAN_CUDA(expr)                                        \
  do {                                                         \
    cudaError_t _e = (expr);                                   \
    if (_e != cudaSuccess) {                                   \
      plan_release(p);                                         \
      return cuda_fail(_e, #expr, __FILE__, __LINE__);         \
    }                                                          \
  } while (0)

  Slot *table = nullptr;
  PLAN_TRY(plan_alloc(&table, (size_t)p->cap + 1, s));
  p->table = table;
  PLAN_TRY(plan_alloc(&p->next, (size_t)p->nb, s));
  PLAN_TRY(plan_alloc(&p->pslot, (size_t)p->np, s));
  PLAN_TRY(plan_alloc(&p->out_off, (size_t)p->np + 1, s));
  Scratch pcount;
  PLAN_TRY(pcount.alloc((size_t)(p->np + 1) * sizeof(uint32_t), s));
  if (join_type == CYL_JOIN_FULL_OUTER) {
    PLAN_TRY(plan_alloc(&p->bsurv, (size_t)p->nb, s));
    PLAN_TRY(plan_alloc(&p->bsurv_pos, (size_t)p->nb + 1, s));
    PLAN_CUDA(cudaMemsetAsync(p->bsurv, 1, (size_t)(p->nb ? p->nb : 1), s));
  }

  HashJoinArgs a;
  fill_args(p, &a);
  a.pcount = pcount.as<uint32_t>();
  {
    ProfScope ps("hash_join_init_slots", s);
    init_slots_kernel<<<grid1d(p->cap + 1), 256, 0, s>>>(table, p->cap + 1);
    note_launch();
  }
  PLAN_CUDA(cudaGetLastError());
  if (p->nb > 0) {
    ProfScope ps("hash_join_build", s);
    hash_build_kernel<<<grid1d(p->nb), 256, 0, s>>>(a);
    note_launch();
    PLAN_CUDA(cudaGetLastError());
  }
  if (p->np > 0) {
    ProfScope ps("hash_join_probe_count", s);
    hash_probe_count_kernel<<<grid1d(p->np), 256, 0, s>>>(a);
    note_launch();
    PLAN_CUDA(cudaGetLastError());
  }
  {
    ProfScope ps("hash_join_scan", s);
    PLAN_TRY(exclusive_scan<uint32_t>(pcount.as<uint32_t>(), p->np, p->out_off, s));
  }
  int64_t totals[2] = {0, 0};
  PLAN_CUDA(cudaMemcpyAsync(&totals[0], p->out_off + p->np, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  if (join_type == CYL_JOIN_FULL_OUTER) {
    PLAN_TRY(exclusive_scan<uint8_t>(p->bsurv, p->nb, p->bsurv_pos, s));
    PLAN_CUDA(cudaMemcpyAsync(&totals[1], p->bsurv_pos + p->nb, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  }
  PLAN_CUDA(cudaStreamSynchronize(s));
  p->m_probe = totals[0];
  p->m_extra = totals[1];
  *num_rows = p->m_probe + p->m_extra;
  *plan = p;
  return CYL_OK;
}

int cyl_join_materialize(cyl_join_plan_t *plan, int64_t *l_idx, int64_t *r_idx, void *stream) {
  if (!plan) return fail(CYL_Invalid, "cyl_join_materialize: null plan");
  cudaStream_t s = (cudaStream_t)stream;
  int rc = plan->algorithm == CYL_JOIN_HASH ? hash_join_materialize(plan, l_idx, r_idx, s) : sort_join_materialize(plan, l_idx, r_idx, s);
  plan->stream = s;
  plan_release(plan);
  return rc;
}

void cyl_join_plan_free(cyl_join_plan_t *plan) { plan_release(plan); }

}  // extern "C"
