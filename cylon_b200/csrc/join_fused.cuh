// join_fused.cuh -- shared declarations of the fused join paths (join_fused.cu generic widths, join_fused8.cu all-8-byte).
#pragma once
#include "common.cuh"

namespace cylb {
constexpr int kFused8MaxCols = 4;
bool fused8_eligible(const cyl_column_t *lcols, int nlcols, const cyl_column_t *rcols, int nrcols);
// build side b*, probe side p*; bout/pout: output column pointers of each side (capacity rows each)
int fused8_join(const cyl_column_t *bcols, int nbc, int bkey, int64_t nb, const cyl_column_t *pcols, int npc, int pkey, int64_t np,
                void *const *bout, void *const *pout, int64_t capacity, int64_t *num_rows, cudaStream_t s);
// Radix partition of an all-8-byte table (<= kFused8MaxCols columns) on mix64(key column) >> shift into `nbins` = 2^(64-shift)
// bins: rec receives the rows as packed records (ncols words each) in bin order, bin_off[nbins + 1] the bin boundaries.
// Shared by the fused join and the fused group-by.
int f8_partition(const cyl_column_t *cols, int ncols, int key, int64_t n, int64_t nbins, int shift, uint64_t *rec, int64_t *bin_off,
                 cudaStream_t s);
// The same join in two steps, so that one table can be partitioned while the other is still on its way (distributed join):
struct F8Partitioned;
void fused8_bins_for(int build_words, int64_t build_rows, int64_t *nbins, int *shift);
int fused8_partition_side(const cyl_column_t *cols, int ncols, int key, int64_t n, int64_t nbins, int shift, F8Partitioned **out, cudaStream_t s);
int fused8_probe(const F8Partitioned *b, const F8Partitioned *p, void *const *bout, void *const *pout, int64_t capacity, int64_t *num_rows,
                 cudaStream_t s);
void fused8_side_free(F8Partitioned *p);
int64_t fused8_side_bins(const F8Partitioned *p);
int fused8_side_shift(const F8Partitioned *p);
// Test hook: force 2^bits bins in fused8_join (0 = choose from the build size). More bins than needed is always valid.
void fused8_force_bins_log2(int bits);
// Measurement knobs (cyl_tune_set): key 1 = tile ranking of the radix passes, key 2 = chain linking of the join.
void fused8_tune(int key, int value);
}  // namespace cylb
