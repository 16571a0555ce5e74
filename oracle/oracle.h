/*
 * oracle.h -- CPU restatement of the reference's hot path. TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build, load or call this. The product (cylon_b200/) never links or calls it.
 *
 * Every function follows the cited reference code (paths relative to
 * /root/reference/cpp/src/cylon/). All pointers are HOST pointers. Parity status: PINNED,
 * two ways. (1) The reference's own golden files and known-answer vectors (tests/golden/,
 * tests/test_oracle_golden.py): murmur3, int/double/string row hashing, partition ids, INNER
 * hash/sort join per rank at world 1/2/4, 2-column-key join, union/subtract/intersect per rank,
 * unique, groupby totals. (2) The reference's own code compiled in place and run as the checker
 * (oracle/_ref/libcylon_ref.so, oracle/ref_shim/, tests/test_oracle_vs_reference.py): its
 * DistributedJoin (all four types, hash and sort, one- and two-column keys), DistributedUnion/
 * Subtract/Intersect, DistributedUnique, Shuffle and DistributedHashGroupBy on 1-4 ranks
 * (threads + an in-memory channel in place of MPI), its aggregation kernels, and null inputs.
 * Still "parity unpinned": string columns in set operations and string join keys -- the
 * reference's cross-table string comparator is broken, so there the oracle
 * restates the intended semantics of the cited code.
 */
#ifndef CYLON_ORACLE_H_
#define CYLON_ORACLE_H_

#include "../include/cylon_b200_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* util/murmur3.cpp:76-117 */
uint32_t orc_murmur3_x86_32(const void *key, int len, uint32_t seed);

/* HashPartitionKernel::UpdateHash, arrow/arrow_partition_kernels.cpp:95-109,151-168,216-240,285-304.
 * In-place: h[i] = 31*h[i] + colhash(v[i]). */
int orc_update_hash(const cyl_column_t *col, uint32_t *h);

/* TableRowIndexHash, arrow/arrow_comparator.cpp:412-440: zero h, UpdateHash for every column. */
int orc_row_hash(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint32_t *h);

/* MapToHashPartitions, partition/partition.cpp:92-166 (+ Partition(), arrow_partition_kernels.cpp:71-93). */
int orc_hash_partition(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint32_t num_partitions,
                       uint32_t *pid, uint64_t *hist);

/* Split, partition/partition.cpp:27-90 + arrow/arrow_kernels.cpp:40-72: stable scatter, expressed as
 * the permutation that lists the rows of partition 0, then 1, ... each in input order. */
int orc_split_indices(const uint32_t *pid, int64_t nrows, uint32_t num_partitions, int64_t *perm,
                      int64_t *part_offsets /* num_partitions+1 */);

/* copy_array_by_indices (numeric / fixed binary), util/copy_arrray.cpp:24-50,87-110:
 * out[i] = idx[i] == -1 ? null : src[idx[i]]; out_validity is a bitmap (may be NULL). */
int orc_gather_fixed(const void *src, int32_t byte_width, const int64_t *idx, int64_t m, void *out,
                     uint8_t *out_validity, int64_t *out_null_count);

/* HashJoin, join/hash_join.cpp:21-90,92-186,188-307. Allocates *l_idx / *r_idx (free with orc_free). */
int orc_hash_join(const cyl_column_t *lkeys, const cyl_column_t *rkeys, int32_t nkeys, int64_t nl,
                  int64_t nr, int32_t join_type, int64_t **l_idx, int64_t **r_idx, int64_t *m);

/* SortJoin, join/sort_join.cpp:396-556 (merge of two argsorted sides; the single-column variants
 * :36-190,:214-370 produce the same pairs). */
int orc_sort_join(const cyl_column_t *lkeys, const cyl_column_t *rkeys, int32_t nkeys, int64_t nl,
                  int64_t nr, int32_t join_type, int64_t **l_idx, int64_t **r_idx, int64_t *m);

void orc_free(void *p);

/* make_groups, groupby/hash_groupby.cpp:92-126: dense group id in first-seen order + first-row mask. */
int orc_groupby_ids(const cyl_column_t *keys, int32_t nkeys, int64_t nrows, int64_t *gid,
                    uint8_t *first_mask, int64_t *num_groups);

/* aggregate<op,T>, groupby/hash_groupby.cpp:142-201 with the kernels of
 * compute/aggregate_kernels.hpp:410-485 (SUM in T, COUNT -> int64, MIN init max(), MAX init
 * numeric_limits<T>::min() -- i.e. DBL_MIN/FLT_MIN for floating point). out has num_groups entries
 * of T (SUM/MIN/MAX) or int64 (COUNT). */
int orc_aggregate(const cyl_column_t *col, const int64_t *gid, int64_t num_groups, int32_t op, void *out);

/* Union/Subtract/Intersect, table.cpp:531-721 (comparator bug E3 of SURVEY.md fixed: binary columns
 * compare the intended tables). mask_l/mask_r: one byte per row, 1 = row is in the output
 * (mask_r only used by UNION). */
int orc_setop_mask(const cyl_column_t *lcols, const cyl_column_t *rcols, int32_t ncols, int64_t nl,
                   int64_t nr, int32_t op, uint8_t *mask_l, uint8_t *mask_r);

/* Unique (first=true), table.cpp:923-982. */
int orc_unique_mask(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint8_t *mask);

/* Order-independent checksum of a table (sum over rows of a 64-bit mix of the row's values; a null
 * contributes 0 bits and validity 0). Not in the reference: it lets full-size results be compared
 * without sorting. Bit-identical to cyl_table_checksum. */
int orc_table_checksum(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint64_t *checksum);

/* Whole local inner/outer join of two (int64 key, double payload) tables, materialised:
 * hash_join.cpp:188-307 + join_utils.cpp:126-181. Used as the timed CPU baseline ("port"). */
int orc_join_i64_f64(const int64_t *lk, const double *lv, int64_t nl, const int64_t *rk, const double *rv,
                     int64_t nr, int32_t join_type, int32_t algorithm, int64_t *m_out, uint64_t *checksum);

#ifdef __cplusplus
}
#endif
#endif
