/*
 * oracle.c -- CPU restatement of the reference's hot path. TEST INFRASTRUCTURE ONLY
 * (see oracle.h for who may call it and for the parity-pinning status).
 *
 * Single-threaded, like one rank of the reference. Reference paths below are relative to
 * /root/reference/cpp/src/cylon/.
 */
#define _GNU_SOURCE
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <limits.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * murmur3 -- util/murmur3.cpp:53-61 (fmix32), :76-117 (MurmurHash3_x86_32)
 * ---------------------------------------------------------------------------------------- */
static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

uint32_t orc_murmur3_x86_32(const void *key, int len, uint32_t seed) {
  const uint8_t *data = (const uint8_t *)key;
  const int nblocks = len / 4;
  uint32_t h1 = seed;
  const uint32_t c1 = 0xcc9e2d51u, c2 = 0x1b873593u;
  for (int i = 0; i < nblocks; i++) {
    uint32_t k1;
    memcpy(&k1, data + 4 * i, 4); /* little-endian block read */
    k1 *= c1;
    k1 = rotl32(k1, 15);
    k1 *= c2;
    h1 ^= k1;
    h1 = rotl32(h1, 13);
    h1 = h1 * 5 + 0xe6546b64u;
  }
  const uint8_t *tail = data + nblocks * 4;
  uint32_t k1 = 0;
  switch (len & 3) {
    case 3: k1 ^= (uint32_t)tail[2] << 16; /* fallthrough */
    case 2: k1 ^= (uint32_t)tail[1] << 8;  /* fallthrough */
    case 1:
      k1 ^= tail[0];
      k1 *= c1;
      k1 = rotl32(k1, 15);
      k1 *= c2;
      h1 ^= k1;
  }
  h1 ^= (uint32_t)len;
  h1 ^= h1 >> 16;
  h1 *= 0x85ebca6bu;
  h1 ^= h1 >> 13;
  h1 *= 0xc2b2ae35u;
  h1 ^= h1 >> 16;
  return h1;
}

/* ------------------------------------------------------------------------------------------
 * per-type value hash: the factory of arrow_partition_kernels.cpp:307-330
 *   ints / temporal  -> ModuloPartitionKernel:   static_cast<uint32_t>(value)       (:84,:104)
 *   float / double   -> NumericHashPartitionKernel: murmur3(&value, sizeof, 0)      (:138-143)
 *   fixed binary     -> murmur3(bytes, byte_width, 0)                               (:208-210)
 *   string / binary  -> murmur3(bytes, len, 0)                                      (:271-275)
 * The validity bitmap is never consulted.
 * ---------------------------------------------------------------------------------------- */
enum { K_UNSUPPORTED = 0, K_MODULO, K_MURMUR_FIXED, K_MURMUR_VAR };

static int hash_kind(int32_t type) {
  switch (type) {
    case CYL_UINT8: case CYL_INT8: case CYL_UINT16: case CYL_INT16: case CYL_UINT32: case CYL_INT32:
    case CYL_UINT64: case CYL_INT64: case CYL_DATE32: case CYL_DATE64: case CYL_TIMESTAMP:
    case CYL_TIME32: case CYL_TIME64:
      return K_MODULO;
    case CYL_FLOAT: case CYL_DOUBLE: case CYL_FIXED_SIZE_BINARY:
      return K_MURMUR_FIXED;
    case CYL_STRING: case CYL_BINARY:
      return K_MURMUR_VAR;
    default:
      return K_UNSUPPORTED;
  }
}

static int is_signed_int(int32_t type) {
  switch (type) {
    case CYL_INT8: case CYL_INT16: case CYL_INT32: case CYL_INT64: case CYL_DATE32: case CYL_DATE64:
    case CYL_TIMESTAMP: case CYL_TIME32: case CYL_TIME64:
      return 1;
    default:
      return 0;
  }
}

/* value of an integer column widened exactly as C++ would before static_cast<uint32_t> */
static inline int64_t load_int(const cyl_column_t *c, int64_t i) {
  const uint8_t *p = (const uint8_t *)c->data + i * c->byte_width;
  if (is_signed_int(c->type)) {
    switch (c->byte_width) {
      case 1: return *(const int8_t *)p;
      case 2: { int16_t v; memcpy(&v, p, 2); return v; }
      case 4: { int32_t v; memcpy(&v, p, 4); return v; }
      default: { int64_t v; memcpy(&v, p, 8); return v; }
    }
  } else {
    switch (c->byte_width) {
      case 1: return *(const uint8_t *)p;
      case 2: { uint16_t v; memcpy(&v, p, 2); return v; }
      case 4: { uint32_t v; memcpy(&v, p, 4); return v; }
      default: { uint64_t v; memcpy(&v, p, 8); return (int64_t)v; }
    }
  }
}

static inline uint32_t value_hash(const cyl_column_t *c, int kind, int64_t i) {
  switch (kind) {
    case K_MODULO:
      return (uint32_t)load_int(c, i);
    case K_MURMUR_FIXED:
      return orc_murmur3_x86_32((const uint8_t *)c->data + i * c->byte_width, c->byte_width, 0);
    default: {
      int32_t b = c->offsets[i], e = c->offsets[i + 1];
      return orc_murmur3_x86_32((const uint8_t *)c->data + b, e - b, 0);
    }
  }
}

int orc_update_hash(const cyl_column_t *col, uint32_t *h) {
  int kind = hash_kind(col->type);
  if (kind == K_UNSUPPORTED) return CYL_ExecutionError;
  for (int64_t i = 0; i < col->length; i++) h[i] = 31u * h[i] + value_hash(col, kind, i);
  return CYL_OK;
}

int orc_row_hash(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint32_t *h) {
  memset(h, 0, (size_t)nrows * sizeof(uint32_t));
  for (int32_t c = 0; c < ncols; c++) {
    if (cols[c].length != nrows) return CYL_Invalid;
    int rc = orc_update_hash(&cols[c], h);
    if (rc) return rc;
  }
  return CYL_OK;
}

/* get_partitioner, arrow_partition_kernels.cpp:51-61 */
static inline uint32_t partition_of(uint32_t h, uint32_t p) { return (p & (p - 1)) == 0 ? (h & (p - 1)) : (h % p); }

int orc_hash_partition(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint32_t num_partitions,
                       uint32_t *pid, uint64_t *hist) {
  if (ncols < 1 || num_partitions == 0) return CYL_Invalid;
  memset(pid, 0, (size_t)nrows * sizeof(uint32_t));
  memset(hist, 0, (size_t)num_partitions * sizeof(uint64_t));
  for (int32_t c = 0; c + 1 < ncols; c++) { /* partition.cpp:146-155 */
    int rc = orc_update_hash(&cols[c], pid);
    if (rc) return rc;
  }
  const cyl_column_t *last = &cols[ncols - 1]; /* partition.cpp:158 */
  int kind = hash_kind(last->type);
  if (kind == K_UNSUPPORTED) return CYL_ExecutionError;
  for (int64_t i = 0; i < nrows; i++) {
    uint32_t h = 31u * pid[i] + value_hash(last, kind, i);
    uint32_t p = partition_of(h, num_partitions);
    pid[i] = p;
    hist[p]++;
  }
  return CYL_OK;
}

int orc_split_indices(const uint32_t *pid, int64_t nrows, uint32_t num_partitions, int64_t *perm,
                      int64_t *part_offsets) {
  int64_t *cursor = (int64_t *)calloc(num_partitions + 1, sizeof(int64_t));
  if (!cursor) return CYL_OutOfMemory;
  for (int64_t i = 0; i < nrows; i++) {
    if (pid[i] >= num_partitions) { free(cursor); return CYL_IndexError; }
    cursor[pid[i] + 1]++;
  }
  for (uint32_t p = 0; p < num_partitions; p++) cursor[p + 1] += cursor[p];
  memcpy(part_offsets, cursor, (num_partitions + 1) * sizeof(int64_t));
  for (int64_t i = 0; i < nrows; i++) perm[cursor[pid[i]]++] = i; /* builders[pid]->UnsafeAppend: input order kept */
  free(cursor);
  return CYL_OK;
}

int orc_gather_fixed(const void *src, int32_t w, const int64_t *idx, int64_t m, void *out, uint8_t *out_validity,
                     int64_t *out_null_count) {
  int64_t nulls = 0;
  if (out_validity) memset(out_validity, 0, (size_t)((m + 7) / 8));
  for (int64_t i = 0; i < m; i++) {
    if (idx[i] == -1) { /* copy_arrray.cpp:39-41: UnsafeAppendNull */
      memset((uint8_t *)out + i * w, 0, (size_t)w);
      nulls++;
    } else {
      memcpy((uint8_t *)out + i * w, (const uint8_t *)src + idx[i] * w, (size_t)w);
      if (out_validity) out_validity[i >> 3] |= (uint8_t)(1u << (i & 7));
    }
  }
  if (out_null_count) *out_null_count = nulls;
  return CYL_OK;
}

void orc_free(void *p) { free(p); }

/* ------------------------------------------------------------------------------------------
 * Row comparison over (up to) two tables -- arrow_comparator.cpp:107-130,205-243 (numeric: by value),
 * :136-181,:246-281 (binary: lexicographic; E3 fixed), :511-543 (column chain).
 * ---------------------------------------------------------------------------------------- */
static inline int cmp_value(const cyl_column_t *a, int64_t i, const cyl_column_t *b, int64_t j) {
  switch (a->type) {
    case CYL_FLOAT: {
      float x, y;
      memcpy(&x, (const uint8_t *)a->data + i * 4, 4);
      memcpy(&y, (const uint8_t *)b->data + j * 4, 4);
      return (x > y) - (x < y);
    }
    case CYL_DOUBLE: {
      double x, y;
      memcpy(&x, (const uint8_t *)a->data + i * 8, 8);
      memcpy(&y, (const uint8_t *)b->data + j * 8, 8);
      return (x > y) - (x < y);
    }
    case CYL_STRING: case CYL_BINARY: {
      int32_t ab = a->offsets[i], al = a->offsets[i + 1] - ab;
      int32_t bb = b->offsets[j], bl = b->offsets[j + 1] - bb;
      int r = memcmp((const uint8_t *)a->data + ab, (const uint8_t *)b->data + bb, (size_t)(al < bl ? al : bl));
      if (r) return (r > 0) - (r < 0);
      return (al > bl) - (al < bl);
    }
    case CYL_FIXED_SIZE_BINARY: {
      int r = memcmp((const uint8_t *)a->data + i * a->byte_width, (const uint8_t *)b->data + j * b->byte_width,
                     (size_t)a->byte_width);
      return (r > 0) - (r < 0);
    }
    case CYL_UINT64: {
      uint64_t x = (uint64_t)load_int(a, i), y = (uint64_t)load_int(b, j);
      return (x > y) - (x < y);
    }
    default: {
      int64_t x = load_int(a, i), y = load_int(b, j);
      return (x > y) - (x < y);
    }
  }
}

static inline int cmp_rows(const cyl_column_t *a, int64_t i, const cyl_column_t *b, int64_t j, int32_t ncols) {
  for (int32_t c = 0; c < ncols; c++) {
    int r = cmp_value(&a[c], i, &b[c], j);
    if (r) return r;
  }
  return 0;
}

/* equal_to of the single-key hash join -- arrow_comparator.cpp:240-243: operator== on the values, so a NaN key equals
 * nothing (cmp_value's three-way result is 0 for NaN, as the reference's subtraction trick :226-238 also yields). */
static inline int eq_single_key(const cyl_column_t *a, int64_t i, const cyl_column_t *b, int64_t j) {
  if (a->type == CYL_DOUBLE) {
    double x, y;
    memcpy(&x, (const uint8_t *)a->data + i * 8, 8);
    memcpy(&y, (const uint8_t *)b->data + j * 8, 8);
    return x == y;
  }
  if (a->type == CYL_FLOAT) {
    float x, y;
    memcpy(&x, (const uint8_t *)a->data + i * 4, 4);
    memcpy(&y, (const uint8_t *)b->data + j * 4, 4);
    return x == y;
  }
  return cmp_value(a, i, b, j) == 0;
}

static int64_t next_pow2(int64_t n) {
  int64_t c = 16;
  while (c < n) c <<= 1;
  return c;
}

/* ------------------------------------------------------------------------------------------
 * growable (l, r) index-pair vector -- the two std::vector<int64_t> of hash_join.cpp / sort_join.cpp
 * ---------------------------------------------------------------------------------------- */
typedef struct { int64_t *l, *r; int64_t n, cap; } pairvec_t;

static int pv_init(pairvec_t *v, int64_t cap) {
  v->n = 0;
  v->cap = cap < 16 ? 16 : cap;
  v->l = (int64_t *)malloc((size_t)v->cap * 8);
  v->r = (int64_t *)malloc((size_t)v->cap * 8);
  return (v->l && v->r) ? 0 : CYL_OutOfMemory;
}
static inline int pv_push(pairvec_t *v, int64_t l, int64_t r) {
  if (v->n == v->cap) {
    v->cap *= 2;
    v->l = (int64_t *)realloc(v->l, (size_t)v->cap * 8);
    v->r = (int64_t *)realloc(v->r, (size_t)v->cap * 8);
    if (!v->l || !v->r) return CYL_OutOfMemory;
  }
  v->l[v->n] = l;
  v->r[v->n] = r;
  v->n++;
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * Hash join -- hash_join.cpp. The unordered_multimap<int64,int64> keyed by row index with
 * (precomputed u32 row hash, row equality) is restated as bucket heads + a next[] chain.
 * ---------------------------------------------------------------------------------------- */
int orc_hash_join(const cyl_column_t *lkeys, const cyl_column_t *rkeys, int32_t nkeys, int64_t nl, int64_t nr,
                  int32_t join_type, int64_t **l_idx, int64_t **r_idx, int64_t *m) {
  for (int32_t c = 0; c < nkeys; c++)
    if (lkeys[c].type != rkeys[c].type) return CYL_Invalid; /* hash_join.cpp:195-197 */
  /* build-side choice, hash_join.cpp:207-232: build_right==1 -> build from right, probe from left */
  int build_right;
  switch (join_type) {
    case CYL_JOIN_LEFT: build_right = 1; break;
    case CYL_JOIN_RIGHT: build_right = 0; break;
    default: build_right = (nl <= nr) ? 0 : 1; break;
  }
  const cyl_column_t *bk = build_right ? rkeys : lkeys, *pk = build_right ? lkeys : rkeys;
  const int64_t nb = build_right ? nr : nl, np = build_right ? nl : nr;

  uint32_t *hb = (uint32_t *)malloc((size_t)(nb + 1) * 4), *hp = (uint32_t *)malloc((size_t)(np + 1) * 4);
  int64_t nbuckets = next_pow2(nb);
  int64_t *head = (int64_t *)malloc((size_t)nbuckets * 8), *next = (int64_t *)malloc((size_t)(nb + 1) * 8);
  uint8_t *b_survivor = NULL, *p_survivor = NULL;
  if (!hb || !hp || !head || !next) return CYL_OutOfMemory;
  int rc;
  if ((rc = orc_row_hash(bk, nkeys, nb, hb)) || (rc = orc_row_hash(pk, nkeys, np, hp))) return rc; /* :239-244 */
  memset(head, 0xff, (size_t)nbuckets * 8);
  /* build, hash_join.cpp:289-291; insert at tail order is irrelevant to the multiset */
  for (int64_t i = nb - 1; i >= 0; i--) {
    int64_t b = hb[i] & (nbuckets - 1);
    next[i] = head[b];
    head[b] = i;
  }
  pairvec_t out;
  if ((rc = pv_init(&out, nb < np ? nb : np))) return rc;
  if (join_type == CYL_JOIN_FULL_OUTER) {
    b_survivor = (uint8_t *)malloc((size_t)nb + 1);
    p_survivor = (uint8_t *)calloc((size_t)np + 1, 1);
    if (!b_survivor || !p_survivor) return CYL_OutOfMemory;
    memset(b_survivor, 1, (size_t)nb + 1);
  }
  const int fast_i64 = (nkeys == 1 && bk[0].type == CYL_INT64);
  const int64_t *bk64 = (const int64_t *)bk[0].data, *pk64 = (const int64_t *)pk[0].data;
  for (int64_t i = 0; i < np; i++) {
    int found = 0;
    for (int64_t j = head[hp[i] & (nbuckets - 1)]; j >= 0; j = next[j]) {
      if (hb[j] != hp[i]) continue;
      int eq = fast_i64 ? (bk64[j] == pk64[i]) : (nkeys == 1 ? eq_single_key(bk, j, pk, i) : (cmp_rows(bk, j, pk, i, nkeys) == 0));
      if (!eq) continue;
      found = 1;
      if (b_survivor) b_survivor[j] = 0;
      rc = build_right ? pv_push(&out, i, j) : pv_push(&out, j, i);
      if (rc) return rc;
    }
    /* probe_hash_map_with_fill :36-54 / _outer :56-90: unmatched probe row -> (-1, i) */
    if (!found && join_type != CYL_JOIN_INNER) {
      if (join_type == CYL_JOIN_FULL_OUTER) { p_survivor[i] = 1; continue; } /* appended after the build survivors, :80-88 */
      rc = build_right ? pv_push(&out, i, -1) : pv_push(&out, -1, i);
      if (rc) return rc;
    }
  }
  if (join_type == CYL_JOIN_FULL_OUTER) {
    for (int64_t j = 0; j < nb; j++)
      if (b_survivor[j]) { rc = build_right ? pv_push(&out, -1, j) : pv_push(&out, j, -1); if (rc) return rc; }
    for (int64_t i = 0; i < np; i++) /* probe_tab_survivors, :82-88 */
      if (p_survivor[i]) { rc = build_right ? pv_push(&out, i, -1) : pv_push(&out, -1, i); if (rc) return rc; }
  }
  free(hb); free(hp); free(head); free(next); free(b_survivor); free(p_survivor);
  *l_idx = out.l;
  *r_idx = out.r;
  *m = out.n;
  return CYL_OK;
}

/* ------------------------------------------------------------------------------------------
 * Sort join -- sort_join.cpp:396-556. argsort each side by the key tuple (SortIndicesMultiColumns,
 * arrow_kernels.cpp:412-457), then merge equal-key groups.
 * ---------------------------------------------------------------------------------------- */
typedef struct { const cyl_column_t *cols; int32_t ncols; } sortctx_t;

static int argsort_cmp(const void *a, const void *b, void *ctx) {
  const sortctx_t *s = (const sortctx_t *)ctx;
  int64_t i = *(const int64_t *)a, j = *(const int64_t *)b;
  int r = cmp_rows(s->cols, i, s->cols, j, s->ncols);
  if (r) return r;
  return (i > j) - (i < j);
}

static int64_t *argsort(const cyl_column_t *cols, int32_t ncols, int64_t n) {
  int64_t *idx = (int64_t *)malloc((size_t)(n + 1) * 8);
  if (!idx) return NULL;
  for (int64_t i = 0; i < n; i++) idx[i] = i;
  sortctx_t ctx = {cols, ncols};
  qsort_r(idx, (size_t)n, 8, argsort_cmp, &ctx);
  return idx;
}

/* advance_multi_index, sort_join.cpp:372-394: [*pos, end) is the next run of equal keys */
static inline int64_t run_end(const cyl_column_t *cols, int32_t ncols, const int64_t *sorted, int64_t n, int64_t pos) {
  int64_t e = pos + 1;
  while (e < n && cmp_rows(cols, sorted[pos], cols, sorted[e], ncols) == 0) e++;
  return e;
}

int orc_sort_join(const cyl_column_t *lkeys, const cyl_column_t *rkeys, int32_t nkeys, int64_t nl, int64_t nr,
                  int32_t join_type, int64_t **l_idx, int64_t **r_idx, int64_t *m) {
  for (int32_t c = 0; c < nkeys; c++)
    if (lkeys[c].type != rkeys[c].type) return CYL_Invalid;
  int64_t *ls = argsort(lkeys, nkeys, nl), *rs = argsort(rkeys, nkeys, nr);
  if (!ls || !rs) return CYL_OutOfMemory;
  const int keep_l = (join_type == CYL_JOIN_LEFT || join_type == CYL_JOIN_FULL_OUTER);
  const int keep_r = (join_type == CYL_JOIN_RIGHT || join_type == CYL_JOIN_FULL_OUTER);
  pairvec_t out;
  int rc;
  if ((rc = pv_init(&out, nl < nr ? nl : nr))) return rc;
  int64_t lp = 0, rp = 0;
  while (lp < nl && rp < nr) {
    int64_t le = run_end(lkeys, nkeys, ls, nl, lp), re = run_end(rkeys, nkeys, rs, nr, rp);
    int com = cmp_rows(lkeys, ls[lp], rkeys, rs[rp], nkeys);
    if (com == 0) { /* :462-476 cross product */
      for (int64_t a = lp; a < le; a++)
        for (int64_t b = rp; b < re; b++)
          if ((rc = pv_push(&out, ls[a], rs[b]))) return rc;
      lp = le;
      rp = re;
    } else if (com < 0) { /* :477-487 */
      if (keep_l)
        for (int64_t a = lp; a < le; a++)
          if ((rc = pv_push(&out, ls[a], -1))) return rc;
      lp = le;
    } else { /* :488-499 */
      if (keep_r)
        for (int64_t b = rp; b < re; b++)
          if ((rc = pv_push(&out, -1, rs[b]))) return rc;
      rp = re;
    }
  }
  if (keep_l) /* :503-512 */
    for (; lp < nl; lp++)
      if ((rc = pv_push(&out, ls[lp], -1))) return rc;
  if (keep_r) /* :514-523 */
    for (; rp < nr; rp++)
      if ((rc = pv_push(&out, -1, rs[rp]))) return rc;
  free(ls);
  free(rs);
  *l_idx = out.l;
  *r_idx = out.r;
  *m = out.n;
  return CYL_OK;
}

/* ------------------------------------------------------------------------------------------
 * Row set over one or two tables: ska::bytell_hash_set<int64, RowIndexHash, RowIndexEqualTo>
 * restated as chained buckets. An entry is (table tag, row); "first inserted row wins".
 * ---------------------------------------------------------------------------------------- */
typedef struct {
  const cyl_column_t *cols[2];
  const uint32_t *hash[2];
  int32_t ncols;
  int64_t nbuckets, n, cap;
  int64_t *head, *next, *entry; /* entry = tag<<62 | row */
} rowset_t;

static int rs_init(rowset_t *s, int64_t expected) {
  s->nbuckets = next_pow2(expected);
  s->cap = expected + 16;
  s->n = 0;
  s->head = (int64_t *)malloc((size_t)s->nbuckets * 8);
  s->next = (int64_t *)malloc((size_t)s->cap * 8);
  s->entry = (int64_t *)malloc((size_t)s->cap * 8);
  if (!s->head || !s->next || !s->entry) return CYL_OutOfMemory;
  memset(s->head, 0xff, (size_t)s->nbuckets * 8);
  return 0;
}
static void rs_free(rowset_t *s) { free(s->head); free(s->next); free(s->entry); }

/* returns the entry slot of an equal row, or -1 */
static inline int64_t rs_find(const rowset_t *s, int tag, int64_t row) {
  uint32_t h = s->hash[tag][row];
  for (int64_t e = s->head[h & (s->nbuckets - 1)]; e >= 0; e = s->next[e]) {
    int t = (int)(s->entry[e] >> 62);
    int64_t r = s->entry[e] & (((int64_t)1 << 62) - 1);
    if (s->hash[t][r] == h && cmp_rows(s->cols[t], r, s->cols[tag], row, s->ncols) == 0) return e;
  }
  return -1;
}
/* returns 1 if inserted (unique so far), 0 if an equal row was already there */
static inline int rs_insert(rowset_t *s, int tag, int64_t row) {
  if (rs_find(s, tag, row) >= 0) return 0;
  uint32_t h = s->hash[tag][row];
  int64_t b = h & (s->nbuckets - 1);
  s->entry[s->n] = ((int64_t)tag << 62) | row;
  s->next[s->n] = s->head[b];
  s->head[b] = s->n;
  s->n++;
  return 1;
}

int orc_groupby_ids(const cyl_column_t *keys, int32_t nkeys, int64_t nrows, int64_t *gid, uint8_t *first_mask,
                    int64_t *num_groups) {
  uint32_t *h = (uint32_t *)malloc((size_t)(nrows + 1) * 4);
  int64_t *entry_gid = (int64_t *)malloc((size_t)(nrows + 16) * 8);
  if (!h || !entry_gid) return CYL_OutOfMemory;
  int rc = orc_row_hash(keys, nkeys, nrows, h);
  if (rc) return rc;
  rowset_t s = {{keys, keys}, {h, h}, nkeys, 0, 0, 0, NULL, NULL, NULL};
  if ((rc = rs_init(&s, nrows))) return rc;
  int64_t unique = 0;
  for (int64_t i = 0; i < nrows; i++) { /* hash_groupby.cpp:113-121 */
    int64_t e = rs_find(&s, 0, i);
    if (e < 0) {
      rs_insert(&s, 0, i);
      entry_gid[s.n - 1] = unique;
      gid[i] = unique++;
      first_mask[i] = 1;
    } else {
      gid[i] = entry_gid[e];
      first_mask[i] = 0;
    }
  }
  *num_groups = unique;
  rs_free(&s);
  free(h);
  free(entry_gid);
  return CYL_OK;
}

#define AGG_LOOP(T, INIT_MIN, INIT_MAX)                                                       \
  do {                                                                                        \
    const T *v = (const T *)col->data;                                                        \
    T *o = (T *)out;                                                                          \
    if (op == CYL_AGG_SUM) {                                                                  \
      for (int64_t g = 0; g < num_groups; g++) o[g] = 0;                                      \
      for (int64_t i = 0; i < n; i++) o[gid[i]] += v[i];                                      \
    } else if (op == CYL_AGG_MIN) {                                                           \
      for (int64_t g = 0; g < num_groups; g++) o[g] = INIT_MIN;                               \
      for (int64_t i = 0; i < n; i++) o[gid[i]] = v[i] < o[gid[i]] ? v[i] : o[gid[i]];        \
    } else {                                                                                  \
      for (int64_t g = 0; g < num_groups; g++) o[g] = INIT_MAX;                               \
      for (int64_t i = 0; i < n; i++) o[gid[i]] = o[gid[i]] < v[i] ? v[i] : o[gid[i]];        \
    }                                                                                         \
  } while (0)

/* MeanKernel / VarianceKernel (VAR, STDDEV), aggregate_kernels.hpp:344-405. States accumulate in the column's own C type
 * (integers wrap, float sums stay float); MEAN divides sum (T) by count (int64) with C++'s usual conversions and stores T;
 * VAR / STDDEV finish in double with ddof = 1, the default of VarKernelOptions (:72), and 0 for a single-row group. */
#define MEANVAR_LOOP(T)                                                                                         \
  do {                                                                                                          \
    const T *v = (const T *)col->data;                                                                          \
    T *sum = (T *)calloc((size_t)num_groups + 1, sizeof(T)), *sq = (T *)calloc((size_t)num_groups + 1, sizeof(T)); \
    int64_t *cnt = (int64_t *)calloc((size_t)num_groups + 1, sizeof(int64_t));                                  \
    if (!sum || !sq || !cnt) return CYL_OutOfMemory;                                                            \
    for (int64_t i = 0; i < n; i++) {                                                                           \
      sq[gid[i]] += v[i] * v[i];                                                                                \
      sum[gid[i]] += v[i];                                                                                      \
      cnt[gid[i]] += 1;                                                                                         \
    }                                                                                                           \
    for (int64_t g = 0; g < num_groups; g++) {                                                                  \
      if (op == ORC_AGG_MEAN) {                                                                                 \
        ((T *)out)[g] = cnt[g] ? (T)(sum[g] / cnt[g]) : (T)0;                                                   \
      } else {                                                                                                  \
        double r = 0;                                                                                           \
        if (cnt[g] > 1) {                                                                                       \
          double mean = (double)sum[g] / (double)cnt[g], mean_sq = (double)sq[g] / (double)cnt[g];              \
          r = (double)cnt[g] * (mean_sq - mean * mean) / (double)(cnt[g] - 1);                                  \
          if (op == ORC_AGG_STDDEV) r = sqrt(r);                                                                \
        }                                                                                                       \
        ((double *)out)[g] = r;                                                                                 \
      }                                                                                                         \
    }                                                                                                           \
    free(sum); free(sq); free(cnt);                                                                             \
  } while (0)

enum { ORC_AGG_MEAN = 4, ORC_AGG_VAR = 5, ORC_AGG_STDDEV = 8 }; /* AggregationOpId, aggregate_kernels.hpp:40-50 */

int orc_aggregate(const cyl_column_t *col, const int64_t *gid, int64_t num_groups, int32_t op, void *out) {
  const int64_t n = col->length;
  if (op == ORC_AGG_MEAN || op == ORC_AGG_VAR || op == ORC_AGG_STDDEV) {
    switch (col->type) {
      case CYL_UINT8: MEANVAR_LOOP(uint8_t); break;
      case CYL_INT8: MEANVAR_LOOP(int8_t); break;
      case CYL_UINT16: MEANVAR_LOOP(uint16_t); break;
      case CYL_INT16: MEANVAR_LOOP(int16_t); break;
      case CYL_UINT32: MEANVAR_LOOP(uint32_t); break;
      case CYL_INT32: MEANVAR_LOOP(int32_t); break;
      case CYL_UINT64: MEANVAR_LOOP(uint64_t); break;
      case CYL_INT64: MEANVAR_LOOP(int64_t); break;
      case CYL_FLOAT: MEANVAR_LOOP(float); break;
      case CYL_DOUBLE: MEANVAR_LOOP(double); break;
      default: return CYL_ExecutionError;
    }
    return CYL_OK;
  }
  if (op == CYL_AGG_COUNT) { /* CountKernel, aggregate_kernels.hpp:432-447: counts every row */
    int64_t *o = (int64_t *)out;
    for (int64_t g = 0; g < num_groups; g++) o[g] = 0;
    for (int64_t i = 0; i < n; i++) o[gid[i]] += 1;
    return CYL_OK;
  }
  if (op != CYL_AGG_SUM && op != CYL_AGG_MIN && op != CYL_AGG_MAX) return CYL_NotImplemented;
  /* MinKernel init numeric_limits<T>::max() (:459), MaxKernel init numeric_limits<T>::min() (:477) */
  switch (col->type) {
    case CYL_UINT8: AGG_LOOP(uint8_t, UINT8_MAX, 0); break;
    case CYL_INT8: AGG_LOOP(int8_t, INT8_MAX, INT8_MIN); break;
    case CYL_UINT16: AGG_LOOP(uint16_t, UINT16_MAX, 0); break;
    case CYL_INT16: AGG_LOOP(int16_t, INT16_MAX, INT16_MIN); break;
    case CYL_UINT32: AGG_LOOP(uint32_t, UINT32_MAX, 0); break;
    case CYL_INT32: AGG_LOOP(int32_t, INT32_MAX, INT32_MIN); break;
    case CYL_UINT64: AGG_LOOP(uint64_t, UINT64_MAX, 0); break;
    case CYL_INT64: AGG_LOOP(int64_t, INT64_MAX, INT64_MIN); break;
    case CYL_FLOAT: AGG_LOOP(float, FLT_MAX, FLT_MIN); break;
    case CYL_DOUBLE: AGG_LOOP(double, DBL_MAX, DBL_MIN); break;
    default: return CYL_ExecutionError; /* hash_groupby.cpp:234-247 */
  }
  return CYL_OK;
}

int orc_setop_mask(const cyl_column_t *lcols, const cyl_column_t *rcols, int32_t ncols, int64_t nl, int64_t nr,
                   int32_t op, uint8_t *mask_l, uint8_t *mask_r) {
  for (int32_t c = 0; c < ncols; c++)
    if (lcols[c].type != rcols[c].type) return CYL_Invalid; /* VerifyTableSchema */
  uint32_t *hl = (uint32_t *)malloc((size_t)(nl + 1) * 4), *hr = (uint32_t *)malloc((size_t)(nr + 1) * 4);
  if (!hl || !hr) return CYL_OutOfMemory;
  int rc;
  if ((rc = orc_row_hash(lcols, ncols, nl, hl)) || (rc = orc_row_hash(rcols, ncols, nr, hr))) return rc;
  rowset_t s = {{lcols, rcols}, {hl, hr}, ncols, 0, 0, 0, NULL, NULL, NULL};
  if ((rc = rs_init(&s, nl + nr))) return rc;
  if (op == CYL_SETOP_UNION) { /* table.cpp:556-586 */
    for (int64_t i = 0; i < nl; i++) mask_l[i] = (uint8_t)rs_insert(&s, 0, i);
    for (int64_t i = 0; i < nr; i++) mask_r[i] = (uint8_t)rs_insert(&s, 1, i);
  } else if (op == CYL_SETOP_SUBTRACT) { /* table.cpp:630-645 */
    for (int64_t i = 0; i < nl; i++) mask_l[i] = (uint8_t)rs_insert(&s, 0, i);
    for (int64_t i = 0; i < nr; i++) {
      int64_t e = rs_find(&s, 1, i);
      if (e >= 0) mask_l[s.entry[e]] = 0;
    }
    if (mask_r) memset(mask_r, 0, (size_t)nr);
  } else if (op == CYL_SETOP_INTERSECT) { /* table.cpp:689-703 */
    for (int64_t i = 0; i < nl; i++) { rs_insert(&s, 0, i); mask_l[i] = 0; }
    for (int64_t i = 0; i < nr; i++) {
      int64_t e = rs_find(&s, 1, i);
      if (e >= 0) mask_l[s.entry[e]] = 1;
    }
    if (mask_r) memset(mask_r, 0, (size_t)nr);
  } else {
    return CYL_Invalid;
  }
  rs_free(&s);
  free(hl);
  free(hr);
  return CYL_OK;
}

int orc_unique_mask(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint8_t *mask) {
  uint32_t *h = (uint32_t *)malloc((size_t)(nrows + 1) * 4);
  if (!h) return CYL_OutOfMemory;
  int rc = orc_row_hash(cols, ncols, nrows, h);
  if (rc) return rc;
  rowset_t s = {{cols, cols}, {h, h}, ncols, 0, 0, 0, NULL, NULL, NULL};
  if ((rc = rs_init(&s, nrows))) return rc;
  for (int64_t i = 0; i < nrows; i++) mask[i] = (uint8_t)rs_insert(&s, 0, i); /* table.cpp:948-951 */
  rs_free(&s);
  free(h);
  return CYL_OK;
}

/* ------------------------------------------------------------------------------------------
 * Order-independent table checksum (not in the reference; it is the "checksum of checksums" used to
 * compare full-size GPU results with the oracle without sorting): sum over rows of a mix of the row's
 * values. Must match cyl_table_checksum in cylon_b200/csrc bit for bit.
 * ---------------------------------------------------------------------------------------- */
static inline uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

int orc_table_checksum(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint64_t *checksum) {
  uint64_t total = 0;
  for (int64_t i = 0; i < nrows; i++) {
    uint64_t acc = 0x9e3779b97f4a7c15ULL;
    for (int32_t c = 0; c < ncols; c++) {
      const cyl_column_t *col = &cols[c];
      int valid = col->validity ? ((col->validity[i >> 3] >> (i & 7)) & 1) : 1;
      uint64_t bits = 0;
      if (valid) {
        if (col->type == CYL_STRING || col->type == CYL_BINARY) {
          int32_t b = col->offsets[i], e = col->offsets[i + 1];
          const uint8_t *p = (const uint8_t *)col->data + b;
          bits = ((uint64_t)orc_murmur3_x86_32(p, e - b, 1) << 32) | orc_murmur3_x86_32(p, e - b, 0);
        } else if (col->byte_width <= 8) {
          memcpy(&bits, (const uint8_t *)col->data + i * col->byte_width, (size_t)col->byte_width);
        } else {
          const uint8_t *p = (const uint8_t *)col->data + i * col->byte_width;
          bits = ((uint64_t)orc_murmur3_x86_32(p, col->byte_width, 1) << 32) | orc_murmur3_x86_32(p, col->byte_width, 0);
        }
      }
      acc = mix64(acc ^ bits) + (uint64_t)valid;
    }
    total += acc;
  }
  *checksum = total;
  return CYL_OK;
}

/* ------------------------------------------------------------------------------------------
 * Timed CPU baseline: cylon::Join on (int64 key, double value) tables, result materialised the way
 * build_final_table does (join_utils.cpp:126-181: all left columns then all right columns).
 * ---------------------------------------------------------------------------------------- */
int orc_join_i64_f64(const int64_t *lk, const double *lv, int64_t nl, const int64_t *rk, const double *rv, int64_t nr,
                     int32_t join_type, int32_t algorithm, int64_t *m_out, uint64_t *checksum) {
  cyl_column_t l = {CYL_INT64, 8, nl, 0, NULL, NULL, lk}, r = {CYL_INT64, 8, nr, 0, NULL, NULL, rk};
  int64_t *li = NULL, *ri = NULL, m = 0;
  int rc = algorithm == CYL_JOIN_HASH ? orc_hash_join(&l, &r, 1, nl, nr, join_type, &li, &ri, &m)
                                      : orc_sort_join(&l, &r, 1, nl, nr, join_type, &li, &ri, &m);
  if (rc) return rc;
  int64_t *olk = (int64_t *)malloc((size_t)(m + 1) * 8), *ork = (int64_t *)malloc((size_t)(m + 1) * 8);
  double *olv = (double *)malloc((size_t)(m + 1) * 8), *orv = (double *)malloc((size_t)(m + 1) * 8);
  uint8_t *vl = (uint8_t *)malloc((size_t)(m / 8 + 1)), *vr = (uint8_t *)malloc((size_t)(m / 8 + 1));
  if (!olk || !ork || !olv || !orv || !vl || !vr) return CYL_OutOfMemory;
  int64_t nulls_l = 0, nulls_r = 0;
  orc_gather_fixed(lk, 8, li, m, olk, vl, &nulls_l);
  orc_gather_fixed(lv, 8, li, m, olv, NULL, NULL);
  orc_gather_fixed(rk, 8, ri, m, ork, vr, &nulls_r);
  orc_gather_fixed(rv, 8, ri, m, orv, NULL, NULL);
  if (checksum) {
    cyl_column_t out[4] = {{CYL_INT64, 8, m, nulls_l, nulls_l ? vl : NULL, NULL, olk},
                           {CYL_DOUBLE, 8, m, nulls_l, nulls_l ? vl : NULL, NULL, olv},
                           {CYL_INT64, 8, m, nulls_r, nulls_r ? vr : NULL, NULL, ork},
                           {CYL_DOUBLE, 8, m, nulls_r, nulls_r ? vr : NULL, NULL, orv}};
    orc_table_checksum(out, 4, m, checksum);
  }
  *m_out = m;
  free(li); free(ri); free(olk); free(ork); free(olv); free(orv); free(vl); free(vr);
  return CYL_OK;
}
