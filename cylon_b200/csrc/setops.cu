// setops.cu -- row-set masks for Union / Subtract / Intersect / Unique, and mask compaction.
//
// Replaces (reference, relative to cpp/src/cylon/): table.cpp:531-603 (Union), :605-663 (Subtract), :665-721
// (Intersect), :923-982 (Unique, first = true) and the arrow::compute::Filter calls that follow them
// (table.cpp:567,586,652,710,968). Semantics kept: outputs are SETS (first occurrence in input order survives),
// Union = distinct left rows ++ right rows not seen before, Intersect/Subtract = distinct left rows that do / do
// not occur on the right. The reference's cross-table string comparator is broken; this
// implements the intended value equality for every column type.
#include "rowtable.cuh"
#include "scan.cuh"

namespace cylb {

static inline int grid1d(int64_t n) {
  int64_t b = div_up(n, 256);
  const int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

struct SetArgs {
  DCols lk, rk;
  RowTable t;
  int64_t nl, nr;
  int32_t exact;
  int32_t op;
  unsigned long long seed;
  const uint32_t *lslot, *rslot;
  uint8_t *mask_l, *mask_r;
  uint8_t *hit;  // [nl] set for left representatives that some right row equals
};

__device__ __forceinline__ bool rep_equals(const SetArgs &a, unsigned long long rep, bool row_is_right, int64_t row) {
  const bool rep_right = (rep & kRightBit) != 0;
  const int64_t r = (int64_t)(rep & ~kRightBit);
  const DCols &ra = rep_right ? a.rk : a.lk;
  const DCols &rb = row_is_right ? a.rk : a.lk;
  return rows_equal(ra, r, rb, row);
}

// mask[i] = row i is the representative (first occurrence) of its slot; hashed mode verifies equality
__global__ void __launch_bounds__(256) set_flag_kernel(const __grid_constant__ SetArgs a, int right_side) {
  const int64_t n = right_side ? a.nr : a.nl;
  const uint32_t *slot = right_side ? a.rslot : a.lslot;
  uint8_t *mask = right_side ? a.mask_r : a.mask_l;
  const unsigned long long bits = right_side ? kRightBit : 0ull;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long rep = a.t.slots[slot[i]].rep;
    const bool mine = rep == ((unsigned long long)i | bits);
    mask[i] = mine ? 1 : 0;
    if (!a.exact && !mine && !rep_equals(a, rep, right_side != 0, i)) a.t.flags[1] = 1;
  }
}

// right rows probe the left-only table: hit[rep] = 1 (table.cpp:637-643, :696-702)
__global__ void __launch_bounds__(256) set_probe_right_kernel(const __grid_constant__ SetArgs a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.nr; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long tag = rt_sanitize(row_tag(a.rk, i, a.exact != 0, a.seed), a.exact != 0);
    const uint32_t s = rt_find(a.t, tag);
    if (s == kNoSlot) continue;
    const unsigned long long rep = a.t.slots[s].rep;
    if (rep == kRtEmpty) continue;  // the reserved slot, never claimed
    if (a.exact || rows_equal(a.lk, (int64_t)rep, a.rk, i)) {
      a.hit[rep] = 1;
    } else {
      a.t.flags[1] = 1;  // same tag, different row: cannot tell whether an equal left row exists -> retry
    }
  }
}

__global__ void __launch_bounds__(256) set_combine_kernel(uint8_t *__restrict__ mask_l, const uint8_t *__restrict__ hit, int64_t nl, int want_hit) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nl; i += (int64_t)gridDim.x * blockDim.x)
    mask_l[i] = (mask_l[i] && (hit[i] != 0) == (want_hit != 0)) ? 1 : 0;
}

__global__ void __launch_bounds__(256) mask_indices_kernel(const uint8_t *__restrict__ mask, const int64_t *__restrict__ pos, int64_t n,
                                                           int64_t *__restrict__ idx) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    if (mask[i]) idx[pos[i]] = i;
}

__global__ void __launch_bounds__(256) gather_validity_kernel(const uint8_t *__restrict__ vsrc, const int64_t *__restrict__ idx, int64_t m,
                                                              uint32_t *__restrict__ vdst) {
  const int64_t mpad = (m + 31) & ~31ll;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < mpad; i += (int64_t)gridDim.x * blockDim.x) {
    bool v = false;
    if (i < m) {
      const int64_t id = idx[i];
      v = id >= 0 && (vsrc == nullptr || bit_is_set(vsrc, id));
    }
    const unsigned bits = __ballot_sync(0xffffffffu, v);
    if (lane_id() == 0) vdst[i >> 5] = bits;
  }
}

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_rowset_mask(const cyl_column_t *lcols, const cyl_column_t *rcols, int32_t ncols, int64_t nl, int64_t nr, int32_t op,
                    uint8_t *mask_l, uint8_t *mask_r, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (ncols < 1) return fail(CYL_Invalid, "cyl_rowset_mask: needs at least one column");
  if (op < CYL_SETOP_UNION || op > CYL_SETOP_UNIQUE) return fail(CYL_Invalid, "cyl_rowset_mask: unknown operation %d", op);
  SetArgs a;
  memset(&a, 0, sizeof(a));
  CYL_TRY(make_dcols(lcols, ncols, nl, &a.lk, "cyl_rowset_mask(left)"));
  if (op == CYL_SETOP_UNIQUE) {
    nr = 0;
    a.rk = a.lk;
  } else {
    for (int c = 0; c < ncols; c++)
      if (lcols[c].type != rcols[c].type)  // VerifyTableSchema, table.cpp:540,615,674
        return fail(CYL_Invalid, "The fields of two tables are not similar (column %d: type %d vs %d)", c, lcols[c].type, rcols[c].type);
    CYL_TRY(make_dcols(rcols, ncols, nr, &a.rk, "cyl_rowset_mask(right)"));
  }
  a.nl = nl;
  a.nr = nr;
  a.exact = is_exact_key(a.lk) ? 1 : 0;
  a.op = op;
  a.mask_l = mask_l;
  a.mask_r = mask_r;
  const bool both_inserted = op == CYL_SETOP_UNION;
  if (!both_inserted && mask_r && nr > 0) CYL_CUDA(cudaMemsetAsync(mask_r, 0, (size_t)nr, s));
  if (nl == 0 && (!both_inserted || nr == 0)) return CYL_OK;

  Scratch lslot, rslot, hit;
  CYL_TRY(lslot.alloc((size_t)(nl + 1) * sizeof(uint32_t), s));
  if (both_inserted) CYL_TRY(rslot.alloc((size_t)(nr + 1) * sizeof(uint32_t), s));
  const bool probes = op == CYL_SETOP_SUBTRACT || op == CYL_SETOP_INTERSECT;
  if (probes) CYL_TRY(hit.alloc((size_t)nl + 1, s));
  a.lslot = lslot.as<uint32_t>();
  a.rslot = rslot.as<uint32_t>();
  a.hit = hit.as<uint8_t>();

  const int64_t inserted = nl + (both_inserted ? nr : 0);
  if (inserted > (1ll << 30)) return fail(CYL_CapacityError, "cyl_rowset_mask: %lld rows to insert (max 2^30 per GPU: 32-bit slot indices)", (long long)inserted);
  const int64_t max_cap = next_pow2(inserted * 2 < 1024 ? 1024 : inserted * 2);
  int64_t cap = rowtable_initial_capacity(inserted);
  RowTableHost rt;
  for (int attempt = 0;; attempt++) {
    if (attempt > 64) return fail(CYL_ExecutionError, "cyl_rowset_mask: could not place all rows");
    CYL_TRY(rt.alloc(cap, s));
    CYL_TRY(rowtable_insert(rt, a.lk, nl, a.exact != 0, a.seed, 0, lslot.as<uint32_t>()));
    if (both_inserted) CYL_TRY(rowtable_insert(rt, a.rk, nr, a.exact != 0, a.seed, kRightBit, rslot.as<uint32_t>()));
    int overflow = 0, collision = 0;
    CYL_TRY(rt.read_flags(&overflow, &collision));
    if (overflow) {
      if (cap >= max_cap) return fail(CYL_ExecutionError, "cyl_rowset_mask: table overflow at full capacity");
      cap = cap * 4 > max_cap ? max_cap : cap * 4;
      continue;
    }
    a.t = rt.t;
    if (nl > 0) {
      set_flag_kernel<<<grid1d(nl), 256, 0, s>>>(a, 0);
      CYL_LAUNCH_CHECK();
    }
    if (both_inserted && nr > 0) {
      set_flag_kernel<<<grid1d(nr), 256, 0, s>>>(a, 1);
      CYL_LAUNCH_CHECK();
    }
    if (probes) {
      CYL_CUDA(cudaMemsetAsync(hit.p, 0, (size_t)nl + 1, s));
      if (nr > 0) {
        set_probe_right_kernel<<<grid1d(nr), 256, 0, s>>>(a);
        CYL_LAUNCH_CHECK();
      }
      if (nl > 0) {
        set_combine_kernel<<<grid1d(nl), 256, 0, s>>>(mask_l, hit.as<uint8_t>(), nl, op == CYL_SETOP_INTERSECT);
        CYL_LAUNCH_CHECK();
      }
    }
    if (!a.exact) {
      CYL_TRY(rt.read_flags(&overflow, &collision));
      if (collision) {
        a.seed = a.seed * 6364136223846793005ULL + 1442695040888963407ULL;
        continue;
      }
    }
    break;
  }
  return CYL_OK;
}

int cyl_mask_to_indices(const uint8_t *mask, int64_t n, int64_t *idx, int64_t *count, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!count) return fail(CYL_Invalid, "cyl_mask_to_indices: null output");
  *count = 0;
  if (n == 0) return CYL_OK;
  Scratch pos;
  CYL_TRY(pos.alloc((size_t)(n + 1) * sizeof(int64_t), s));
  CYL_TRY(exclusive_scan<uint8_t>(mask, n, pos.as<int64_t>(), s));
  mask_indices_kernel<<<grid1d(n), 256, 0, s>>>(mask, pos.as<int64_t>(), n, idx);
  CYL_LAUNCH_CHECK();
  int64_t c = 0;
  CYL_CUDA(cudaMemcpyAsync(&c, pos.as<int64_t>() + n, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *count = c;
  return CYL_OK;
}

int cyl_gather_validity(const uint8_t *vsrc, const int64_t *idx, int64_t m, uint8_t *vdst, void *stream) {
  if (m <= 0) return CYL_OK;
  gather_validity_kernel<<<grid1d(m), 256, 0, (cudaStream_t)stream>>>(vsrc, idx, m, reinterpret_cast<uint32_t *>(vdst));
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

}  // extern "C"

// ---- concatenation / validity transport helpers (arrow::ConcatenateTables + CombineChunks, table.cpp:119-128,594-600) ----
namespace cylb {

__global__ void concat_offsets_kernel(const int32_t *__restrict__ a, int64_t na, const int32_t *__restrict__ b, int64_t nb,
                                      int32_t *__restrict__ out) {
  const int32_t a0 = na > 0 ? a[0] : 0;
  const int32_t abytes = na > 0 ? a[na] - a0 : 0;
  const int32_t b0 = nb > 0 ? b[0] : 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= na + nb; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = i <= na ? (na > 0 ? a[i] - a0 : 0) : abytes + (b[i - na] - b0);
}

__global__ void concat_validity_kernel(const uint8_t *__restrict__ va, int64_t na, const uint8_t *__restrict__ vb, int64_t nb,
                                       uint32_t *__restrict__ out) {
  const int64_t n = na + nb, npad = (n + 31) & ~31ll;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npad; i += (int64_t)gridDim.x * blockDim.x) {
    bool v = false;
    if (i < na) v = va ? bit_is_set(va, i) : true;
    else if (i < n) v = vb ? bit_is_set(vb, i - na) : true;
    const unsigned bits = __ballot_sync(0xffffffffu, v);
    if (lane_id() == 0) out[i >> 5] = bits;
  }
}

__global__ void unpack_validity_kernel(const uint8_t *__restrict__ bitmap, int64_t n, uint8_t *__restrict__ bytes) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    bytes[i] = bitmap ? (bit_is_set(bitmap, i) ? 1 : 0) : 1;
}

__global__ void pack_validity_kernel(const uint8_t *__restrict__ bytes, int64_t n, uint32_t *__restrict__ bitmap) {
  const int64_t npad = (n + 31) & ~31ll;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npad; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned bits = __ballot_sync(0xffffffffu, i < n && bytes[i] != 0);
    if (lane_id() == 0) bitmap[i >> 5] = bits;
  }
}

__global__ void offsets_to_lengths_kernel(const int32_t *__restrict__ off, int64_t n, int32_t *__restrict__ len) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) len[i] = off[i + 1] - off[i];
}

__global__ void gather_offsets_at_kernel(const int32_t *__restrict__ off, const int64_t *__restrict__ rows, int64_t n, int32_t *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = off[rows[i]];
}

__global__ void narrow_i64_i32_kernel(const int64_t *__restrict__ in, int64_t n1, int32_t *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n1; i += (int64_t)gridDim.x * blockDim.x) out[i] = (int32_t)in[i];
}

}  // namespace cylb

extern "C" {

int cyl_concat_offsets(const int32_t *a_offsets, int64_t na, const int32_t *b_offsets, int64_t nb, int32_t *out_offsets, void *stream) {
  concat_offsets_kernel<<<grid1d(na + nb + 1), 256, 0, (cudaStream_t)stream>>>(a_offsets, na, b_offsets, nb, out_offsets);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_concat_validity(const uint8_t *va, int64_t na, const uint8_t *vb, int64_t nb, uint8_t *out, void *stream) {
  if (na + nb <= 0) return CYL_OK;
  concat_validity_kernel<<<grid1d(na + nb), 256, 0, (cudaStream_t)stream>>>(va, na, vb, nb, reinterpret_cast<uint32_t *>(out));
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_unpack_validity(const uint8_t *bitmap, int64_t n, uint8_t *bytes, void *stream) {
  if (n <= 0) return CYL_OK;
  unpack_validity_kernel<<<grid1d(n), 256, 0, (cudaStream_t)stream>>>(bitmap, n, bytes);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_pack_validity(const uint8_t *bytes, int64_t n, uint8_t *bitmap, void *stream) {
  if (n <= 0) return CYL_OK;
  pack_validity_kernel<<<grid1d(n), 256, 0, (cudaStream_t)stream>>>(bytes, n, reinterpret_cast<uint32_t *>(bitmap));
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_offsets_to_lengths(const int32_t *offsets, int64_t n, int32_t *lengths, void *stream) {
  if (n <= 0) return CYL_OK;
  offsets_to_lengths_kernel<<<grid1d(n), 256, 0, (cudaStream_t)stream>>>(offsets, n, lengths);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_gather_offsets_at(const int32_t *offsets, const int64_t *rows, int64_t n, int32_t *out, void *stream) {
  if (n <= 0) return CYL_OK;
  gather_offsets_at_kernel<<<grid1d(n), 256, 0, (cudaStream_t)stream>>>(offsets, rows, n, out);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_lengths_to_offsets(const int32_t *lengths, int64_t n, int32_t *offsets, int64_t *total_bytes, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  Scratch o64;
  CYL_TRY(o64.alloc((size_t)(n + 1) * sizeof(int64_t), s));
  CYL_TRY(exclusive_scan<int32_t>(lengths, n, o64.as<int64_t>(), s));
  narrow_i64_i32_kernel<<<grid1d(n + 1), 256, 0, s>>>(o64.as<int64_t>(), n + 1, offsets);
  CYL_LAUNCH_CHECK();
  if (total_bytes) {
    int64_t t = 0;
    CYL_CUDA(cudaMemcpyAsync(&t, o64.as<int64_t>() + n, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    CYL_CUDA(cudaStreamSynchronize(s));
    if (t > 0x7FFFFFFFll) return fail(CYL_CapacityError, "variable-width column needs %lld bytes (> 2^31-1, int32 offsets)", (long long)t);
    *total_bytes = t;
  }
  return CYL_OK;
}

}  // extern "C"
