/*
 * cylon_b200.h -- C-ABI of libcylon_b200.so: the B200 (sm_100a) kernels behind Cylon's distributed
 * relational hot path (hash-partition -> shuffle -> local join | group-by | set operation).
 *
 * Conventions
 *   - plain C, no C++/torch types. Every pointer is a DEVICE pointer unless a comment says "host".
 *   - sizes and row indices are int64_t. A table may hold at most 2^32-2 rows per GPU.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream). Calls are
 *     asynchronous on that stream unless they return a host value (documented per call).
 *   - every call returns a cylon::Code (cyl_code_t; 0 = OK), the convention of cpp/src/cylon/status.hpp:21-63.
 *     cyl_last_error() gives the message of the calling thread's last failure.
 *   - scratch memory comes from the library's caching allocator (cyl_pool_*) and is handed back before the call
 *     returns; it is reused only by later work on the same stream.
 *   - citations are relative to /root/reference/cpp/src/cylon/ and name the reference code each entry
 *     point replaces.
 */
#ifndef CYLON_B200_H_
#define CYLON_B200_H_

#include "cylon_b200_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- runtime ---------------------- */
int cyl_version(void);
const char *cyl_last_error(void); /* host string, valid until the thread's next failing call */
/* Selects `device` for the calling thread and creates its CUDA context.
 * (The reference's counterpart is CylonContext::Init, ctx/cylon_context.cpp:24-30 -- host only.) */
int cyl_init(int32_t device);
int cyl_device_count(int32_t *count /* host */);
int cyl_malloc(void **ptr /* host out */, int64_t bytes);
int cyl_free(void *ptr);
int cyl_memset(void *ptr, int32_t value, int64_t bytes, void *stream);
int cyl_memcpy_h2d(void *dst, const void *src_host, int64_t bytes, void *stream);
int cyl_memcpy_d2h(void *dst_host, const void *src, int64_t bytes, void *stream);
int cyl_memcpy_d2d(void *dst, const void *src, int64_t bytes, void *stream);
int cyl_stream_sync(void *stream);
int cyl_host_alloc_pinned(void **ptr /* host out */, int64_t bytes);
int cyl_host_free_pinned(void *ptr);
/* The library's caching device allocator (every scratch buffer, join plan and table column comes from it): blocks are
 * parked per (device, stream, size class) when released and only reused for work on the same stream, so after warm-up
 * no operator reaches cudaMalloc/cudaFree. The reference's counterpart is the Arrow memory pool proxy
 * (ctx/arrow_memory_pool_utils.hpp:24-62). cyl_pool_trim returns every parked block to the driver. */
int cyl_pool_alloc(void **ptr /* host out */, int64_t bytes, void *stream);
int cyl_pool_free(void *ptr, void *stream);
int cyl_pool_trim(void);
/* Measurement hooks (no counterpart in the reference, which logs std::chrono stage times, e.g. hash_join.cpp:301-304).
 * cyl_launch_count: kernels launched by this library since load. cyl_profile_enable(1) brackets the library's main
 * kernels with CUDA events on their stream; cyl_profile_report synchronises the device and writes
 * "kernel,launches,total_ms" lines (then clears the records). */
long long cyl_launch_count(void);
int cyl_profile_enable(int32_t on);
int cyl_profile_report(char *buf /* host */, int64_t buf_len);

/* ---------------------------------------------------------------- A: row hashing / partition --- */
/* h[i] = chain over `cols` of 31*h + colhash(v), h0 = 0: ints/temporal (uint32_t)v, float/double/fixed
 * binary/string murmur3_x86_32(bytes, seed 0). Validity ignored.
 * Replaces HashPartitionKernel::UpdateHash (arrow/arrow_partition_kernels.cpp:95-109,151-168,216-240,
 * 285-304) and TableRowIndexHash (arrow/arrow_comparator.cpp:396-440). */
int cyl_row_hash(const cyl_column_t *cols /* host array */, int32_t ncols, int64_t nrows, uint32_t *h, void *stream);

/* pid[i] = part(row_hash(i)), part = h & (P-1) if P is a power of two else h % P; hist[p] = #rows.
 * pid may be NULL (histogram only). hist is uint64[P], overwritten.
 * Replaces MapToHashPartitions (partition/partition.cpp:92-166) + HashPartitionKernel::Partition
 * (arrow/arrow_partition_kernels.cpp:71-93,124-149,189-214,258-283). */
int cyl_hash_partition(const cyl_column_t *cols /* host array */, int32_t ncols, int64_t nrows, uint32_t num_partitions,
                       uint32_t *pid, uint64_t *hist, void *stream);

/* ---------------------------------------------------------------- B: split (stable scatter) ---- */
/* dest[i] = part_offsets[pid[i]] + #{j < i : pid[j] == pid[i]}  (row order inside a partition = input order),
 * part_offsets = exclusive scan of the histogram, int64[P+1]. num_partitions <= 1024. Every pid[i] must be below
 * num_partitions; a row whose id is not is counted in no partition and its dest[i] is left untouched (the reference
 * would index past its builder vector there, arrow_kernels.cpp:52).
 * Replaces the row routing of ArrowArraySplitKernel::Split (arrow/arrow_kernels.cpp:40-72). */
int cyl_scatter_map(const uint32_t *pid, int64_t nrows, uint32_t num_partitions, int64_t *dest, int64_t *part_offsets,
                    void *stream);
/* cyl_scatter_map for callers that are not allowed to allocate device memory. (No counterpart in the reference.)
 * Who that is: an engine that hands the whole GPU to its own arena at start-up, so a later cudaMalloc either fails or
 * breaks its memory accounting; or a latency-bound service whose policy forbids allocator calls on the request path
 * because they can stall behind other work on the device. cyl_scatter_map takes its scratch (per-tile histograms,
 * their scanned cursors, the scan's tile sums) from the stream-ordered pool; this variant keeps the same scratch in a
 * fixed 3.0 MB array that is part of the library image, so it calls no cudaMalloc/cudaFree/cudaMallocAsync (or any
 * other allocating routine) at all. Inputs and outputs are still the caller's buffers, and the results are identical
 * to cyl_scatter_map's.
 * Limits and rules:
 *   - tiles x partitions must fit the array: ceil(nrows / tile_rows) * num_partitions <= 262144 counters, where
 *     tile_rows = 1024 for num_partitions <= 128 (and grows with it). That is about 33.5 million rows at 8
 *     partitions, 4.1 million at 64. Larger inputs are rejected with CYL_CapacityError and a message saying so.
 *   - the array exists once per GPU that loaded the library and is held until the process exits.
 *   - the first call on each device looks up the array's device address (with that device current); make that first
 *     call before starting any stream capture.
 *   - all calls on a device share the array. A call overwrites every word it later reads, so stale contents from an
 *     earlier call are harmless, but two calls that can run on the GPU at the same time (different streams or host
 *     threads) race on it: serialise them. */
int cyl_scatter_map_static(const uint32_t *pid, int64_t nrows, uint32_t num_partitions, int64_t *dest, int64_t *part_offsets,
                           void *stream);
/* Same, with pid computed on the fly from the key columns (fuses cyl_hash_partition + cyl_scatter_map;
 * pid_out may be NULL). */
int cyl_partition_map(const cyl_column_t *cols /* host array */, int32_t ncols, int64_t nrows, uint32_t num_partitions,
                      uint32_t *pid_out, int64_t *dest, int64_t *part_offsets, void *stream);
/* dst[dest[i]] = src[i] for a fixed-width column (numeric: arrow_kernels.cpp:33-72, fixed binary :74-113). */
int cyl_scatter_fixed(const void *src, int32_t byte_width, const int64_t *dest, int64_t nrows, void *dst, void *stream);
/* Variable-width column (arrow_kernels.cpp:115-157): dst_offsets (int32[nrows+1], starts at 0) and dst_data
 * (same byte count as the source) receive the rows in dest order. */
int cyl_scatter_varwidth(const int32_t *src_offsets, const uint8_t *src_data, const int64_t *dest, int64_t nrows,
                         int32_t *dst_offsets, uint8_t *dst_data, void *stream);
/* Validity bitmap: bit dest[i] of vdst = bit i of vsrc. vdst has ceil(nrows/32)*4 bytes, overwritten.
 * (The reference drops validity in Split -- SURVEY.md B2; this library keeps it.) */
int cyl_scatter_validity(const uint8_t *vsrc, const int64_t *dest, int64_t nrows, uint8_t *vdst, void *stream);
/* One fused pass for a table whose columns are all fixed width (1/2/4/8/16 bytes): partition ids from the key
 * columns, stable destinations and the scatter of every column, without materialising pid or dest.
 * dst[c] must hold nrows values. part_offsets: int64[P+1] (device). key_col_idx: host array of indices into cols. */
int cyl_partition_scatter_fixed(const cyl_column_t *cols /* host array */, int32_t ncols, const int32_t *key_col_idx /* host */,
                                int32_t nkeys, int64_t nrows, uint32_t num_partitions, void *const *dst /* host array of device ptrs */,
                                int64_t *part_offsets, void *stream);

/* Two-step form of the fused partition + scatter, for the shuffle that writes straight into the receivers' buffers:
 * step 1 counts (tile histograms + scan, kept in the plan) and reports rows per partition on the host (synchronises);
 * the caller exchanges the counts, learns where each partition has to go, and step 2 scatters partition p's rows, in
 * stable order, to dst[p * ncols + c] + (dst_row_base[p] + k) * width for the k-th row of the partition. The
 * destination pointers may be device memory of ANOTHER GPU mapped into this process (cudaIpcOpenMemHandle): the
 * stores then travel over NVLink and no send buffer or separate copy exists. All columns fixed width (1/2/4/8/16).
 * Replaces Split + ArrowAllToAll's per-buffer sends (partition/partition.cpp:27-90, arrow/arrow_all_to_all.cpp:68-151). */
typedef struct cyl_partition_plan cyl_partition_plan_t;
int cyl_partition_plan_create(const cyl_column_t *cols /* host array */, int32_t ncols, const int32_t *key_col_idx /* host */, int32_t nkeys,
                              int64_t nrows, uint32_t num_partitions, cyl_partition_plan_t **plan /* host out */,
                              int64_t *part_rows /* host out [P] */, void *stream);
int cyl_partition_plan_scatter(cyl_partition_plan_t *plan, void *const *dst /* host array [P * ncols] of device pointers */,
                               const int64_t *dst_row_base /* host [P] */, void *stream);
void cyl_partition_plan_free(cyl_partition_plan_t *plan);

/* ---------------------------------------------------------------- scan ------------------------- */
/* out[i] = sum_{j<i} in[j]; out has n+1 entries (out[n] = total). In place (out == in) is NOT allowed. */
int cyl_exclusive_scan_i64(const int64_t *in, int64_t n, int64_t *out, void *stream);

/* ---------------------------------------------------------------- D: joins --------------------- */
typedef struct cyl_join_plan cyl_join_plan_t; /* opaque, host object owning device state */

/* Pass 1 of the hash join: build an open-addressing table in HBM over the build side (LEFT -> right table,
 * RIGHT -> left table, INNER/FULL_OUTER -> the smaller, ties left: join/hash_join.cpp:207-232), probe it and
 * count the output rows. *num_rows (host) receives M; the call synchronises the stream.
 * Replaces ArrayIndexHashJoin (join/hash_join.cpp:188-307) and multi_index_hash_join (:92-186). */
int cyl_hash_join_count(const cyl_column_t *lkeys /* host array */, const cyl_column_t *rkeys /* host array */, int32_t nkeys,
                        int64_t nl, int64_t nr, int32_t join_type, cyl_join_plan_t **plan /* host out */,
                        int64_t *num_rows /* host out */, void *stream);
/* Pass 1 of the sort join: LSD radix argsort of both sides on the key tuple, then per-left-row match ranges
 * in the sorted right side. Replaces SortJoin (join/sort_join.cpp:36-190,214-370,396-556) and its sorters
 * (arrow/arrow_kernels.cpp:294-339,386-457; util/sort.hpp:126-145). Never mutates the key columns (D6). */
int cyl_sort_join_count(const cyl_column_t *lkeys /* host array */, const cyl_column_t *rkeys /* host array */, int32_t nkeys,
                        int64_t nl, int64_t nr, int32_t join_type, cyl_join_plan_t **plan /* host out */,
                        int64_t *num_rows /* host out */, void *stream);
/* Pass 2 (either algorithm): write the M (left row, right row) index pairs, -1 = no partner
 * (probe_hash_map_no_fill/_with_fill/_outer, join/hash_join.cpp:21-90). Order is unspecified. Frees the plan. */
int cyl_join_materialize(cyl_join_plan_t *plan, int64_t *l_idx, int64_t *r_idx, void *stream);
void cyl_join_plan_free(cyl_join_plan_t *plan);

/* Fused fast path of the INNER hash join: radix-partition both tables on the key hash, then build each partition's
 * hash table in SHARED memory and write the joined rows directly (no index vectors, no gather pass). Requirements
 * (cyl_hash_join_fused_supported returns 1): INNER, one key column per side of the same fixed-width type (<= 8 bytes),
 * every column fixed width (1/2/4/8 bytes), at most 8 columns per side. Validity bitmaps of the inputs are ignored, as the
 * reference's join ignores them (keys compare by value slot, copy_arrray.cpp:47 drops source validity): the result has none.
 * out_cols: host array of nlcols + nrcols device pointers (left columns first, the layout build_final_table produces,
 * join/join_utils.cpp:126-181), each with room for `capacity` values. The result size is only known after probing:
 * *num_rows (host out; the call synchronises) always receives the exact count; the rows were written only if
 * *num_rows <= capacity, otherwise repeat the call with capacity >= *num_rows. Row order is unspecified.
 * Replaces ArrayIndexHashJoin + build_final_table + copy_array_by_indices (join/hash_join.cpp:188-307,
 * join/join_utils.cpp:126-181, util/copy_arrray.cpp:24-50). */
int cyl_hash_join_fused_supported(const cyl_column_t *lcols /* host array */, int32_t nlcols, int32_t lkey, const cyl_column_t *rcols /* host array */,
                                  int32_t nrcols, int32_t rkey, int32_t join_type);
int cyl_hash_join_fused_inner(const cyl_column_t *lcols /* host array */, int32_t nlcols, int32_t lkey, const cyl_column_t *rcols /* host array */,
                              int32_t nrcols, int32_t rkey, int64_t nl, int64_t nr, int64_t capacity, void *const *out_cols /* host array */,
                              int64_t *num_rows /* host out */, void *stream);
/* The INNER fused join of all-8-byte tables in two steps: partition each table into the join's bins (any stream), then probe.
 * A distributed join partitions the table that has arrived while the other one is still being shuffled. `like` (may be NULL)
 * makes the second table use the first one's bins; without it the bins are sized for max(nrows, build_rows_hint) build rows.
 * The probe treats `left` as the build side; out_cols / capacity / num_rows as in cyl_hash_join_fused_inner. A side may be
 * probed more than once (capacity retry) and must be freed. Same reference code replaced as cyl_hash_join_fused_inner. */
typedef struct cyl_join_side cyl_join_side_t;
int cyl_hash_join_fused_partition(const cyl_column_t *cols /* host array */, int32_t ncols, int32_t key, int64_t nrows, const cyl_join_side_t *like,
                                  int64_t build_rows_hint, cyl_join_side_t **side /* host out */, void *stream);
int cyl_hash_join_fused_probe(const cyl_join_side_t *left, const cyl_join_side_t *right, int64_t capacity, void *const *out_cols /* host array */,
                              int64_t *num_rows /* host out */, void *stream);
void cyl_hash_join_fused_side_free(cyl_join_side_t *side);
/* The same engine for LEFT / RIGHT / FULL_OUTER joins, keys of one or two columns and tables that mix 4- and 8-byte
 * numeric / temporal columns (at most 4 record words per side; a two-column key uses one of them). out_cols: nlcols +
 * nrcols device buffers of `capacity` values of each column's width, left columns first. left_valid / right_valid: device
 * byte arrays of `capacity` entries, 1 where that side of the output row is present and 0 where it is null (required
 * for the side(s) the join type can leave without a partner, may be NULL otherwise); pack them with cyl_pack_validity.
 * Null slots hold zeros. Source validity is ignored as in the reference's join. Same capacity protocol as above. Used for
 * JoinAlgorithm HASH and SORT alike: both produce this row multiset (join/join.cpp:80-100).
 * Replaces the probe functions and build-side choice of join/hash_join.cpp:21-90,207-232, multi_index_hash_join
 * (:92-186), do_multi_index_sorted_join (join/sort_join.cpp:396-556) and build_final_table (join/join_utils.cpp:126-181). */
int cyl_join_fused_supported(const cyl_column_t *lcols /* host array */, int32_t nlcols, const int32_t *lkeys, const cyl_column_t *rcols /* host array */,
                             int32_t nrcols, const int32_t *rkeys, int32_t nkeys, int32_t join_type);
int cyl_join_fused(const cyl_column_t *lcols /* host array */, int32_t nlcols, const int32_t *lkeys, const cyl_column_t *rcols /* host array */,
                   int32_t nrcols, const int32_t *rkeys, int32_t nkeys, int32_t join_type, int64_t capacity, void *const *out_cols /* host array */,
                   uint8_t *left_valid, uint8_t *right_valid, int64_t *num_rows /* host out */, void *stream);
/* Tuning/test hook with no reference counterpart: force 2^bits radix bins (1..18) in the all-8-byte fused join instead of
 * the count chosen from the build size; 0 restores the automatic choice. More bins than needed is always valid (they
 * are just emptier), so tests use it to reach the many-bin code paths (512-way digits, split histogram) with small
 * tables. Process-wide; do not change it while a join is running. Returns CYL_Invalid outside 0..18. */
int cyl_hash_join_fused_set_bins_log2(int32_t bits);
/* Measurement knob with no reference counterpart: lets two equivalent implementations of an inner step be timed against
 * each other on the same input (bench.py reads CYLON_TUNE=key=value,...). Results never depend on it. No keys are live in
 * this build: the alternatives measured in round 2 lost and were removed (profiles/r02_join_experiments.md). Process-wide. */
int cyl_tune_set(int32_t key, int32_t value);

/* ---------------------------------------------------------------- D4: gather ------------------- */
/* dst[i] = idx[i] < 0 ? 0 : src[idx[i]]; dst_validity (may be NULL) gets bit i = (idx[i] >= 0), it has
 * ceil(m/32)*4 bytes. Source validity is not propagated, as in the reference (copy_arrray.cpp:47): cylon::Join builds
 * its result with validity = (index >= 0) only; the Filter-like callers (set operators, Unique) gather the source bitmap
 * separately with cyl_gather_validity.
 * Replaces copy_array_by_indices (util/copy_arrray.cpp:24-50,87-110,142-292). */
int cyl_gather_fixed(const void *src, int32_t byte_width, const int64_t *idx, int64_t m, void *dst, uint8_t *dst_validity,
                     void *stream);
/* Variable width (util/copy_arrray.cpp:52-85), two steps: offsets (returns total bytes on the host; synchronises),
 * then bytes into a buffer of that size. */
int cyl_gather_varwidth_offsets(const int32_t *src_offsets, const int64_t *idx, int64_t m, int32_t *dst_offsets,
                                int64_t *total_bytes /* host out */, void *stream);
int cyl_gather_varwidth_data(const int32_t *src_offsets, const uint8_t *src_data, const int64_t *idx, int64_t m,
                             const int32_t *dst_offsets, uint8_t *dst_data, void *stream);
/* validity bitmap of a gather by idx: bit i = (idx[i] >= 0); ceil(m/32)*4 bytes (copy_arrray.cpp:39-41,67-68). */
int cyl_index_validity(const int64_t *idx, int64_t m, uint8_t *validity, void *stream);
/* number of negative entries (= nulls a gather by idx produces). Host out; synchronises. */
int cyl_count_negative(const int64_t *idx, int64_t m, int64_t *count /* host out */, void *stream);

/* validity of a filter/take that keeps the source's nulls: bit i = idx[i] >= 0 && (vsrc == NULL || bit idx[i] of vsrc).
 * (arrow::compute::Filter as called at table.cpp:567,586,652,710,968 keeps nulls.) */
int cyl_gather_validity(const uint8_t *vsrc, const int64_t *idx, int64_t m, uint8_t *vdst, void *stream);

/* ---------------------------------------------------------------- G: group-by ------------------ */
/* gid[i] = dense id of row i's key tuple, numbered in first-seen order; first_mask[i] = 1 for the first row of each
 * group (one byte per row); *num_groups is a host out (the call synchronises).
 * Replaces make_groups (groupby/hash_groupby.cpp:92-126). */
int cyl_groupby_hash(const cyl_column_t *keys /* host array */, int32_t nkeys, int64_t nrows, int64_t *gid, uint8_t *first_mask,
                     int64_t *num_groups /* host out */, void *stream);
/* out[g] = op over { col[i] : gid[i] == g }. out holds num_groups values of the column's type (SUM/MIN/MAX) or int64
 * (COUNT). SUM wraps in the column's C type, MIN starts at numeric_limits<T>::max(), MAX at numeric_limits<T>::min()
 * (DBL_MIN/FLT_MIN for floating point -- the reference's behaviour). MEAN: sum (in the column's type) / count with C++'s
 * conversions, stored in the column's type; VAR / STDDEV: double, from sum and sum of squares accumulated in the
 * column's type, ddof = 1 (VarKernelOptions' default), 0 for single-row groups. col is a host pointer to one descriptor.
 * Replaces aggregate<op,T> (groupby/hash_groupby.cpp:142-201) + compute/aggregate_kernels.hpp:344-405,410-485. */
int cyl_aggregate(const cyl_column_t *col /* host */, const int64_t *gid, int64_t num_groups, int32_t op, void *out, void *stream);
/* Fused fast path of HashGroupBy for ONE 8-byte key column (int64 / uint64 / double / date64 / timestamp / time64) and 8-byte
 * value columns (int64 / uint64 / double), all without validity bitmaps, at most 3 distinct value columns and 4 aggregates:
 * make_groups + every aggregate in one sweep (rows radix-partitioned on the key hash, one CTA per bin with its group table
 * in shared memory). agg_value[a] indexes `values`, agg_op[a] is a cyl_agg_op_t. out_key and out_aggs[a] are device
 * buffers of `capacity` 8-byte values (COUNT writes int64). *num_groups (host out; the call synchronises) always
 * receives the exact number of groups; the outputs were written only if it is <= capacity, otherwise repeat the call
 * with capacity >= *num_groups. Group order is unspecified (the reference numbers groups in first-seen order); SUM of
 * doubles adds in an unspecified order (within the 1e-12 relative tolerance of BASELINE.json, not bit for bit).
 * Replaces make_groups + aggregate<op,T> (groupby/hash_groupby.cpp:92-201) + compute/aggregate_kernels.hpp:410-485. */
int cyl_groupby_fused_supported(const cyl_column_t *key /* host */, const cyl_column_t *values /* host array */, int32_t nvalues,
                                const int32_t *agg_value, const int32_t *agg_op, int32_t naggs);
int cyl_groupby_fused(const cyl_column_t *key /* host */, const cyl_column_t *values /* host array */, int32_t nvalues, const int32_t *agg_value,
                      const int32_t *agg_op, int32_t naggs, int64_t capacity, void *out_key, void *const *out_aggs /* host array */,
                      int64_t *num_groups /* host out */, void *stream);

/* ---------------------------------------------------------------- E: set operations / unique --- */
/* One byte per row: mask_l[i] / mask_r[j] = 1 when the row belongs to the result.
 *   UNION:     first occurrence of every distinct left row, plus first occurrence of right rows not on the left
 *   SUBTRACT:  first occurrence of distinct left rows that do not occur on the right   (mask_r zeroed)
 *   INTERSECT: first occurrence of distinct left rows that also occur on the right     (mask_r zeroed)
 *   UNIQUE:    first occurrence of every distinct key tuple of `lcols` (rcols/nr/mask_r ignored)
 * Row equality = per-column value equality over the given columns.
 * Replaces Union/Subtract/Intersect/Unique (table.cpp:531-721,923-982) up to their Filter calls. */
int cyl_rowset_mask(const cyl_column_t *lcols /* host array */, const cyl_column_t *rcols /* host array */, int32_t ncols, int64_t nl,
                    int64_t nr, int32_t op, uint8_t *mask_l, uint8_t *mask_r, void *stream);
/* idx = ascending positions of the non-zero bytes of mask (a stream compaction; feed it to the gather calls).
 * idx must hold n entries; *count is a host out (synchronises). Replaces arrow::compute::Filter's selection. */
int cyl_mask_to_indices(const uint8_t *mask, int64_t n, int64_t *idx, int64_t *count /* host out */, void *stream);

/* ---------------------------------------------------------------- concatenation / transport ---- */
/* Pieces of arrow::ConcatenateTables + CombineChunks (table.cpp:119-128,594-600) for device columns; fixed-width
 * data and variable-width bytes are plain device copies, these do the parts that are not. */
/* out_offsets[0..na+nb] for the concatenation of two variable-width columns (each rebased to start at 0). */
int cyl_concat_offsets(const int32_t *a_offsets, int64_t na, const int32_t *b_offsets, int64_t nb, int32_t *out_offsets, void *stream);
/* bitmap of na+nb bits = va's bits then vb's bits; a NULL input means "all valid". ceil((na+nb)/32)*4 bytes. */
int cyl_concat_validity(const uint8_t *va, int64_t na, const uint8_t *vb, int64_t nb, uint8_t *out, void *stream);
/* validity bitmap <-> one byte per row (the shuffle moves validity as bytes because partition boundaries are not
 * byte aligned in a bitmap; arrow_all_to_all.cpp:97-107 ships the raw bitmap of already-split arrays). */
/* out[i] = offsets[rows[i]] (device arrays): the byte offsets of a variable-width column at n chosen rows, e.g. at the
 * partition boundaries, fetched with one kernel and one copy instead of one 4-byte copy each. Replaces the per-partition
 * value_offset() reads of the string splitter (arrow/arrow_kernels.cpp:115-157). */
int cyl_gather_offsets_at(const int32_t *offsets, const int64_t *rows, int64_t n, int32_t *out, void *stream);
int cyl_unpack_validity(const uint8_t *bitmap /* NULL = all valid */, int64_t n, uint8_t *bytes, void *stream);
int cyl_pack_validity(const uint8_t *bytes, int64_t n, uint8_t *bitmap, void *stream);
/* int32 offsets <-> per-row lengths (the shuffle ships lengths and rebuilds offsets on the receiver).
 * cyl_lengths_to_offsets: *total_bytes is a host out (may be NULL; when given the call synchronises). */
int cyl_offsets_to_lengths(const int32_t *offsets, int64_t n, int32_t *lengths, void *stream);
int cyl_lengths_to_offsets(const int32_t *lengths, int64_t n, int32_t *offsets, int64_t *total_bytes /* host out */, void *stream);

/* ---------------------------------------------------------------- C: shuffle plan -------------- */
/* Host-only arithmetic of the all-to-all (all pointers are HOST pointers; no GPU needed): counts[src * world + dst] =
 * units (rows or bytes) rank src sends to rank dst, known to every rank after the count exchange. For `rank`: where
 * each peer's slice starts / how long it is in the partition-ordered send buffer (send_off, send_cnt) and in the
 * receive buffer, whose chunks are ordered by source rank (recv_off, recv_cnt); *total_recv = rows/bytes received.
 * Replaces the per-buffer header/FIN handshake and the receive-side allocation of
 * arrow/arrow_all_to_all.cpp:68-211 + net/ops/all_to_all.cpp:78-137 with offsets known before any payload moves. */
int cyl_shuffle_plan(const int64_t *counts /* host */, int32_t world, int32_t rank, int64_t *send_off, int64_t *send_cnt, int64_t *recv_off,
                     int64_t *recv_cnt, int64_t *total_recv);

/* ---------------------------------------------------------------- D: sort --------------------- */
/* Row order of a table sorted by `nkeys` columns (keys[0] most significant; ascending[c] != 0: ascending): out_idx[i] = the row
 * that comes i-th. Stable LSD radix sort of order-preserving 64-bit keys, one column per round from the last to the first.
 * Sort columns: numeric / temporal values of at most 8 bytes (else CYL_NotImplemented). Validity is not looked at, as in the
 * reference; floats order by IEEE total order (-0.0 before +0.0, NaN after +inf). Rows that tie on every column keep their input
 * order (the reference's std::sort leaves them unspecified). Replaces SortIndices / SortIndicesMultiColumns
 * (arrow/arrow_kernels.cpp:294-303,412-457) under Sort (table.cpp:291-328, util/arrow_utils.cpp:30-118); the Take that follows
 * is cyl_gather_*. Synchronises. */
int cyl_sort_indices(const cyl_column_t *keys /* host array */, int32_t nkeys, const int32_t *ascending /* host array */, int64_t nrows,
                     int64_t *out_idx /* device, nrows */, void *stream);
/* The range partitioner of DistributedSort: RangePartitionKernel (arrow/arrow_partition_kernels.cpp:334-455) under
 * MapToSortPartitions (partition/partition.cpp:168-198), in three calls with the cross-rank reductions between them
 * (compute::MinMax and mpi::AllReduce in the reference, :384-407):
 *   1. cyl_range_sample_minmax: min and max of this rank's sample as ORDER KEYS (64-bit values whose unsigned order is the value
 *      order, so every rank reduces them with plain unsigned min / max) and the number of sampled non-NaN values. The sample is
 *      rows floor(i * nrows / num_samples), i < num_samples, or every row when num_samples >= nrows (the reference samples at
 *      random unless num_samples == nrows).
 *   2. cyl_range_sample_histogram: the sample counted into num_bins + 2 bins by get_bin_pos (:436-440), evaluated in the
 *      column's own C++ type as the reference's template does (host out, to be summed over the ranks).
 *   3. cyl_range_partition: target partition of EVERY row, bin_to_partition[get_bin_pos(value)] (num_partitions - 1 - that
 *      when descending), and the rows per partition. total_samples <= 0 (nothing sampled on any rank) sends every row to bin 0.
 * All three synchronise. Between 2 and 3, cyl_range_quantile_split (host-only arithmetic, no GPU needed) turns the summed histogram
 * into bin_to_partition as build_bin_to_partition does (:409-427, float_t prefix frequencies against k / num_partitions). */
int cyl_range_quantile_split(const uint64_t *global_counts /* host, num_bins + 2 */, uint32_t num_bins, uint64_t total_samples, uint32_t num_partitions,
                             uint32_t *bin_to_partition /* host out, num_bins + 2 */);
int cyl_range_sample_minmax(const cyl_column_t *col, int64_t nrows, int64_t num_samples, uint64_t *min_key /* host out */,
                            uint64_t *max_key /* host out */, int64_t *count /* host out */, void *stream);
int cyl_range_sample_histogram(const cyl_column_t *col, int64_t nrows, int64_t num_samples, uint64_t min_key, uint64_t max_key, uint32_t num_bins,
                               uint64_t *hist /* host out, num_bins + 2 */, void *stream);
int cyl_range_partition(const cyl_column_t *col, int64_t nrows, uint64_t min_key, uint64_t max_key, int64_t total_samples, uint32_t num_bins,
                        const uint32_t *bin_to_partition /* host array, num_bins + 2 */, uint32_t num_partitions, int32_t ascending,
                        uint32_t *pid /* device out, nrows */, uint64_t *partition_hist /* device out, num_partitions; may be NULL */, void *stream);

/* ---------------------------------------------------------------- checks ----------------------- */
/* Order-independent checksum of a table: sum over rows of a 64-bit mix of the row's values (nulls count as
 * zero bits + validity 0). No counterpart in the reference; it is how full-size results are compared with the
 * CPU oracle without sorting. *checksum is a host out; synchronises. */
int cyl_table_checksum(const cyl_column_t *cols /* host array */, int32_t ncols, int64_t nrows, uint64_t *checksum /* host out */,
                       void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CYLON_B200_H_ */
