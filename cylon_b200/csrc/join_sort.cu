// join_sort.cu -- LSD radix argsort of 64-bit key tags and the sort-merge join built on it.
//
// Replaces (reference, relative to cpp/src/cylon/): join/sort_join.cpp:36-190 (single non-numeric key),
// :214-370 (single numeric key, in place) and :396-556 (multi-column keys), with their sorters
// arrow/arrow_kernels.cpp:294-339,386-457 and util/sort.hpp:126-145 (introsort). The reference argsorts both
// sides with std::sort/introsort through a comparator chain and merges equal-key runs.
//
// Here each side is reduced to (tag, row id) pairs -- tag = the key's bits for a single fixed-width column,
// else a 64-bit hash of the key tuple -- and sorted with a stable 8-bit LSD radix sort (passes whose digit is
// constant over the whole column are skipped). Both sides use the same tag function, so equal keys meet in
// equal-tag runs; when the tag is a hash, candidate pairs inside a run are verified with the reference's value
// equality, which makes the result exact whatever the hash does. The merge is a per-row range search of the
// sorted other side (lower/upper bound on tags), count -> scan -> materialize. The caller's key columns are
// never modified (the reference's in-place variant sorts the caller's buffer).
#include "join_plan.cuh"
#include "scan.cuh"

namespace cylb {

int radix_pass_pairs(const uint64_t *keys_in, const uint32_t *vals_in, uint64_t *keys_out, uint32_t *vals_out, int64_t n, int shift,
                     cudaStream_t stream);

static inline int grid1d(int64_t n) {
  int64_t b = div_up(n, 256);
  const int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

__global__ void __launch_bounds__(256) make_tags_kernel(const __grid_constant__ DCols k, int64_t n, int32_t exact, uint64_t *__restrict__ tags,
                                                        uint32_t *__restrict__ ids) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    // Hashed tags only have to bring equal keys together (candidates are verified by value), so 40 bits are plenty:
    // the low 3 bytes are cleared, their digits are constant and the radix sort skips those passes (5 instead of 8).
    const uint64_t t = row_tag(k, i, exact != 0, 0);
    tags[i] = exact ? t : (t & 0xFFFFFFFFFF000000ULL);
    ids[i] = (uint32_t)i;
  }
}

// 8 digit histograms in one read of the tags: tells the host which radix passes can be skipped
__global__ void __launch_bounds__(256) digit_hist_kernel(const uint64_t *__restrict__ tags, int64_t n, unsigned long long *__restrict__ hist) {
  __shared__ uint32_t sh[8 * 256];
  for (int t = threadIdx.x; t < 8 * 256; t += blockDim.x) sh[t] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t v = tags[i];
#pragma unroll
    for (int d = 0; d < 8; d++) atomicAdd(&sh[d * 256 + ((v >> (8 * d)) & 0xFF)], 1u);
  }
  __syncthreads();
  for (int t = threadIdx.x; t < 8 * 256; t += blockDim.x)
    if (sh[t]) atomicAdd(&hist[t], (unsigned long long)sh[t]);
}

// Sorts (tags, ids) ascending by tag (stable). On return *tags_io / *ids_io point at the sorted buffers (either
// the given ones or the temporaries). Synchronises the stream once (digit histogram readback).
int radix_sort_pairs(uint64_t **tags_io, uint32_t **ids_io, uint64_t **tags_tmp, uint32_t **ids_tmp, int64_t n, cudaStream_t s) {
  if (n <= 1) return CYL_OK;
  Scratch hist;
  CYL_TRY(hist.alloc(8 * 256 * sizeof(unsigned long long), s));
  CYL_CUDA(cudaMemsetAsync(hist.p, 0, 8 * 256 * sizeof(unsigned long long), s));
  int g = grid1d(n);
  if (g > num_sms() * 4) g = num_sms() * 4;
  digit_hist_kernel<<<g, 256, 0, s>>>(*tags_io, n, hist.as<unsigned long long>());
  CYL_LAUNCH_CHECK();
  static thread_local unsigned long long h[8 * 256];
  CYL_CUDA(cudaMemcpyAsync(h, hist.p, sizeof(h), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  for (int d = 0; d < 8; d++) {
    bool constant = false;
    for (int b = 0; b < 256; b++)
      if (h[d * 256 + b] == (unsigned long long)n) constant = true;
    if (constant) continue;
    CYL_TRY(radix_pass_pairs(*tags_io, *ids_io, *tags_tmp, *ids_tmp, n, 8 * d, s));
    uint64_t *tt = *tags_io;
    *tags_io = *tags_tmp;
    *tags_tmp = tt;
    uint32_t *ti = *ids_io;
    *ids_io = *ids_tmp;
    *ids_tmp = ti;
  }
  return CYL_OK;
}

struct SortJoinArgs {
  DCols lk, rk;
  const uint64_t *ltag, *rtag;  // sorted
  const uint32_t *lperm, *rperm;
  int64_t nl, nr;
  int32_t exact;
  int32_t nan_type;  // exact mode with a FLOAT/DOUBLE key: its type id (a NaN key matches nothing), else 0
  int32_t keep_left, keep_right;
  uint32_t *lo, *hi;
  uint32_t *lcount;
  uint8_t *rsurv;
  const int64_t *out_off;
  const int64_t *rsurv_pos;
  int64_t extra_base;
  int64_t *l_idx, *r_idx;
};

__device__ __forceinline__ int64_t lower_bound_u64(const uint64_t *a, int64_t n, uint64_t v) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (a[mid] < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ int64_t upper_bound_from(const uint64_t *a, int64_t from, int64_t n, uint64_t v) {
  // equal-tag runs are short: gallop from the lower bound
  int64_t step = 1, lo = from, hi = from;
  while (hi < n && a[hi] == v) {
    lo = hi + 1;
    hi += step;
    step <<= 1;
  }
  if (hi > n) hi = n;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (a[mid] <= v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// for every sorted-left position: the equal-tag range in the sorted right side and the verified match count
__global__ void __launch_bounds__(256) merge_count_left_kernel(const __grid_constant__ SortJoinArgs a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.nl; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t t = a.ltag[i];
    const bool nan = a.nan_type && key_bits_is_nan(a.nan_type, t);
    const int64_t lo = nan ? 0 : lower_bound_u64(a.rtag, a.nr, t);
    const int64_t hi = nan ? 0 : upper_bound_from(a.rtag, lo, a.nr, t);
    uint32_t c = 0;
    if (a.exact) {
      c = (uint32_t)(hi - lo);
    } else {
      const int64_t li = a.lperm[i];
      for (int64_t q = lo; q < hi; q++) c += rows_equal(a.lk, li, a.rk, a.rperm[q]) ? 1u : 0u;
    }
    a.lo[i] = (uint32_t)lo;
    a.hi[i] = (uint32_t)hi;
    a.lcount[i] = c ? c : (a.keep_left ? 1u : 0u);
  }
}

// RIGHT / FULL_OUTER: a right row survives (gets a null left partner) when no left row equals it
__global__ void __launch_bounds__(256) merge_mark_right_kernel(const __grid_constant__ SortJoinArgs a) {
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < a.nr; q += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t t = a.rtag[q];
    const int64_t lo = lower_bound_u64(a.ltag, a.nl, t);
    bool matched = false;
    if (a.nan_type && key_bits_is_nan(a.nan_type, t)) {
      matched = false;
    } else if (a.exact) {
      matched = lo < a.nl && a.ltag[lo] == t;
    } else {
      const int64_t ri = a.rperm[q];
      for (int64_t i = lo; i < a.nl && a.ltag[i] == t && !matched; i++) matched = rows_equal(a.lk, a.lperm[i], a.rk, ri);
    }
    a.rsurv[q] = matched ? 0 : 1;
  }
}

__global__ void __launch_bounds__(256) merge_write_left_kernel(const __grid_constant__ SortJoinArgs a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.nl; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t o = a.out_off[i];
    const int64_t end = a.out_off[i + 1];
    const int64_t li = a.lperm[i];
    const uint32_t lo = a.lo[i], hi = a.hi[i];
    int64_t written = 0;
    for (uint32_t q = lo; q < hi; q++) {
      const int64_t ri = a.rperm[q];
      if (a.exact || rows_equal(a.lk, li, a.rk, ri)) {
        a.l_idx[o] = li;
        a.r_idx[o] = ri;
        o++;
        written++;
      }
    }
    if (written == 0 && end > o) {  // keep_left
      a.l_idx[o] = li;
      a.r_idx[o] = -1;
    }
  }
}

__global__ void __launch_bounds__(256) merge_write_right_survivors_kernel(const __grid_constant__ SortJoinArgs a) {
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < a.nr; q += (int64_t)gridDim.x * blockDim.x) {
    if (a.rsurv[q]) {
      const int64_t o = a.extra_base + a.rsurv_pos[q];
      a.l_idx[o] = -1;
      a.r_idx[o] = a.rperm[q];
    }
  }
}

static void fill_sort_args(const cyl_join_plan *p, SortJoinArgs *a) {
  memset(a, 0, sizeof(*a));
  a->lk = p->pk;  // sort join: pk = left keys, bk = right keys
  a->rk = p->bk;
  a->lperm = p->lperm;
  a->rperm = p->rperm;
  a->nl = p->nl;
  a->nr = p->nr;
  a->exact = p->exact;
  a->nan_type = (p->exact && (p->pk.c[0].type == CYL_FLOAT || p->pk.c[0].type == CYL_DOUBLE)) ? p->pk.c[0].type : 0;
  a->keep_left = p->join_type == CYL_JOIN_LEFT || p->join_type == CYL_JOIN_FULL_OUTER;
  a->keep_right = p->join_type == CYL_JOIN_RIGHT || p->join_type == CYL_JOIN_FULL_OUTER;
  a->lo = p->lo;
  a->hi = p->hi;
  a->rsurv = p->rsurv;
  a->out_off = p->out_off;
  a->rsurv_pos = p->rsurv_pos;
}

int sort_join_materialize(cyl_join_plan *p, int64_t *l_idx, int64_t *r_idx, cudaStream_t s) {
  SortJoinArgs a;
  fill_sort_args(p, &a);
  a.l_idx = l_idx;
  a.r_idx = r_idx;
  a.extra_base = p->m_probe;
  if (p->nl > 0) {
    merge_write_left_kernel<<<grid1d(p->nl), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  if (p->m_extra > 0) {
    merge_write_right_survivors_kernel<<<grid1d(p->nr), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  return CYL_OK;
}

// argsort one side by tag; returns the sorted tags (scratch owned by caller) and the permutation (plan-owned)
static int sort_side(const DCols &k, int64_t n, bool exact, Scratch *tags_a, Scratch *tags_b, uint32_t **perm_out, uint64_t **tags_out,
                     cudaStream_t s) {
  uint32_t *ids_a = nullptr, *ids_b = nullptr;
  CYL_TRY(tags_a->alloc((size_t)(n + 1) * sizeof(uint64_t), s));
  CYL_TRY(tags_b->alloc((size_t)(n + 1) * sizeof(uint64_t), s));
  CYL_TRY(plan_alloc(&ids_a, (size_t)n + 1, s));
  int rc = plan_alloc(&ids_b, (size_t)n + 1, s);
  if (rc != CYL_OK) {
    pool_free(ids_a, s);
    return rc;
  }
  uint64_t *ta = tags_a->as<uint64_t>(), *tb = tags_b->as<uint64_t>();
  if (n > 0) {
    make_tags_kernel<<<grid1d(n), 256, 0, s>>>(k, n, exact ? 1 : 0, ta, ids_a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) rc = cuda_fail(e, "make_tags_kernel", __FILE__, __LINE__);
  }
  if (rc == CYL_OK) rc = radix_sort_pairs(&ta, &ids_a, &tb, &ids_b, n, s);
  pool_free(ids_b, s);
  if (rc != CYL_OK) {
    pool_free(ids_a, s);
    return rc;
  }
  *perm_out = ids_a;
  *tags_out = ta;
  return CYL_OK;
}

}  // namespace cylb

using namespace cylb;

extern "C" int cyl_sort_join_count(const cyl_column_t *lkeys, const cyl_column_t *rkeys, int32_t nkeys, int64_t nl, int64_t nr,
                                   int32_t join_type, cyl_join_plan_t **plan, int64_t *num_rows, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!plan || !num_rows) return fail(CYL_Invalid, "cyl_sort_join_count: null output");
  *plan = nullptr;
  if (nkeys < 1) return fail(CYL_Invalid, "cyl_sort_join_count: needs at least one key column");
  if (join_type < CYL_JOIN_INNER || join_type > CYL_JOIN_FULL_OUTER) return fail(CYL_Invalid, "cyl_sort_join_count: join type %d", join_type);
  for (int c = 0; c < nkeys; c++)
    if (lkeys[c].type != rkeys[c].type)
      return fail(CYL_Invalid, "left and right index array types are not equal (key %d: %d vs %d)", c, lkeys[c].type, rkeys[c].type);
  DCols lk, rk;
  CYL_TRY(make_dcols(lkeys, nkeys, nl, &lk, "cyl_sort_join_count(left)"));
  CYL_TRY(make_dcols(rkeys, nkeys, nr, &rk, "cyl_sort_join_count(right)"));

  cyl_join_plan *p = new cyl_join_plan();
  p->algorithm = CYL_JOIN_SORT;
  p->join_type = join_type;
  p->stream = s;
  p->nl = nl;
  p->nr = nr;
  p->pk = lk;
  p->bk = rk;
  p->np = nl;
  p->nb = nr;
  p->exact = is_exact_key(lk);

#define SPLAN_TRY(expr)    \
  do {                     \
    int _rc = (expr);      \
    if (_rc != CYL_OK) {   \
      plan_release(p);     \
      return _rc;          \
    }                      \
  } while (0)
#define SPLAN_CUDA(expr)                                   \
  do {                                                     \
    cudaError_t _e = (expr);                               \
    if (_e != cudaSuccess) {                               \
      plan_release(p);                                     \
      return cuda_fail(_e, #expr, __FILE__, __LINE__);     \
    }                                                      \
  } while (0)

  Scratch lt_a, lt_b, rt_a, rt_b, lcount;
  uint64_t *ltag = nullptr, *rtag = nullptr;
  SPLAN_TRY(sort_side(lk, nl, p->exact, &lt_a, &lt_b, &p->lperm, &ltag, s));
  SPLAN_TRY(sort_side(rk, nr, p->exact, &rt_a, &rt_b, &p->rperm, &rtag, s));
  SPLAN_TRY(plan_alloc(&p->lo, (size_t)nl + 1, s));
  SPLAN_TRY(plan_alloc(&p->hi, (size_t)nl + 1, s));
  SPLAN_TRY(plan_alloc(&p->out_off, (size_t)nl + 1, s));
  SPLAN_TRY(lcount.alloc((size_t)(nl + 1) * sizeof(uint32_t), s));

  SortJoinArgs a;
  fill_sort_args(p, &a);
  a.ltag = ltag;
  a.rtag = rtag;
  a.lcount = lcount.as<uint32_t>();
  if (nl > 0) {
    merge_count_left_kernel<<<grid1d(nl), 256, 0, s>>>(a);
    SPLAN_CUDA(cudaGetLastError());
  }
  SPLAN_TRY(exclusive_scan<uint32_t>(lcount.as<uint32_t>(), nl, p->out_off, s));
  int64_t totals[2] = {0, 0};
  SPLAN_CUDA(cudaMemcpyAsync(&totals[0], p->out_off + nl, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  if (a.keep_right) {
    SPLAN_TRY(plan_alloc(&p->rsurv, (size_t)nr + 1, s));
    SPLAN_TRY(plan_alloc(&p->rsurv_pos, (size_t)nr + 1, s));
    a.rsurv = p->rsurv;
    if (nr > 0) {
      merge_mark_right_kernel<<<grid1d(nr), 256, 0, s>>>(a);
      SPLAN_CUDA(cudaGetLastError());
    }
    SPLAN_TRY(exclusive_scan<uint8_t>(p->rsurv, nr, p->rsurv_pos, s));
    SPLAN_CUDA(cudaMemcpyAsync(&totals[1], p->rsurv_pos + nr, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  }
  SPLAN_CUDA(cudaStreamSynchronize(s));
  p->m_probe = totals[0];
  p->m_extra = totals[1];
  *num_rows = p->m_probe + p->m_extra;
  *plan = p;
  return CYL_OK;
}
