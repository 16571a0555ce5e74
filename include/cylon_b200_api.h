/*
 * cylon_b200_api.h -- handle-based C facade over the C++ host API (include/cylon/), for bindings.
 *
 * The reference exposes its operators to non-C++ callers through a facade that names tables by id
 * (cpp/src/cylon/table_api.hpp:37-150, used by java/src/main/native/src/Table.cpp:39,60 and mirrored by
 * python/pycylon/data/table.pyx:314-380). This is the same facade with opaque handles instead of string ids:
 * tables live in HBM, operators run on the context's CUDA stream, and the distributed variants shuffle over NCCL.
 * Every function returns a cylon::Code (0 = OK); cyl_last_error() (cylon_b200.h) gives the message.
 * Host pointers only, except where a cyl_column_t is documented as carrying device pointers.
 */
#ifndef CYLON_B200_API_H_
#define CYLON_B200_API_H_

#include "cylon_b200_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cylon_context cylon_context_t; /* shared_ptr<cylon::CylonContext> */
typedef struct cylon_table cylon_table_t;     /* shared_ptr<cylon::Table> */

/* CylonContext::Init / InitDistributed (ctx/cylon_context.hpp:62-70). */
int cylon_context_init(int32_t device, cylon_context_t **ctx);
int cylon_nccl_unique_id(char id[128]); /* rank 0 creates, every rank passes the same bytes below */
int cylon_context_init_distributed(int32_t rank, int32_t world_size, int32_t device, const char id[128], cylon_context_t **ctx);
/* Ranks as host threads of one process (cylon::net::ThreadConfig): every thread calls this with the same group name and
 * world size and its own rank, then uses its context from that thread only. For tests of the distributed operators on a
 * box with fewer GPUs than ranks; no reference counterpart (the reference's ranks are MPI processes). */
int cylon_context_init_threads(const char *group, int32_t rank, int32_t world_size, int32_t device, cylon_context_t **ctx);
int cylon_context_set_stream(cylon_context_t *ctx, void *cuda_stream);
/* CylonContext::AddConfig (ctx/cylon_context.hpp:85). Keys read by this library: "cylon.join.fused" ("0" = general hash
 * join only), "cylon.shuffle.peer" ("0" = NCCL send/recv instead of NVLink peer stores; must be the same on every rank),
 * "cylon.groupby.fused" / "cylon.setops.fused" ("0" = global row table only), "cylon.groupby.preaggregate" ("0" = raw rows are
 * shuffled when COUNT is among the aggregates, as the reference does), "cylon.join.overlap" ("1" = DistributedJoin partitions the
 * arrived left table on the context's second stream while the right table shuffles; same value on every rank). */
int cylon_context_add_config(cylon_context_t *ctx, const char *key, const char *value);
int cylon_context_rank(const cylon_context_t *ctx);
int cylon_context_world_size(const cylon_context_t *ctx);
int cylon_context_barrier(cylon_context_t *ctx);
int cylon_context_finalize(cylon_context_t *ctx); /* also frees the handle */

/* Table::FromColumns over host Arrow-layout buffers: copies them to HBM (H2D on the context's stream).
 * names may be NULL. (table.hpp:84,100; table_api.cpp FromArrowTable) */
int cylon_table_from_host(cylon_context_t *ctx, const cyl_column_t *host_columns, const char *const *names, int32_t ncols,
                          cylon_table_t **out);
/* Same, over buffers that are already in device memory; the table does not own or copy them. */
int cylon_table_from_device(cylon_context_t *ctx, const cyl_column_t *device_columns, const char *const *names, int32_t ncols,
                            cylon_table_t **out);
int64_t cylon_table_rows(const cylon_table_t *t);
int32_t cylon_table_columns(const cylon_table_t *t);
const char *cylon_table_column_name(const cylon_table_t *t, int32_t col);
/* descriptor with DEVICE pointers (valid while the table lives); *data_bytes = bytes of the data buffer */
int cylon_table_column(const cylon_table_t *t, int32_t col, cyl_column_t *desc, int64_t *data_bytes);
/* device -> host copy of one column into caller buffers (any may be NULL); synchronises */
int cylon_table_column_to_host(const cylon_table_t *t, int32_t col, void *data, int32_t *offsets, uint8_t *validity);
int cylon_table_retain_memory(cylon_table_t *t, int32_t retain); /* Table::retainMemory, table.hpp:168 */
int cylon_table_checksum(const cylon_table_t *t, uint64_t *checksum);
void cylon_table_free(cylon_table_t *t);

/* Join / DistributedJoin (table.hpp:258,269). join_type: cyl_join_type_t, algorithm: cyl_join_algorithm_t. */
int cylon_join(cylon_table_t *left, cylon_table_t *right, int32_t join_type, int32_t algorithm, const int32_t *left_cols,
               const int32_t *right_cols, int32_t nkeys, const char *left_prefix, const char *right_prefix, int32_t distributed,
               cylon_table_t **out);
/* Union/Subtract/Intersect and their Distributed* forms (table.hpp:279-329). op: cyl_setop_t (not UNIQUE). */
int cylon_set_operation(cylon_table_t *left, cylon_table_t *right, int32_t op, int32_t distributed, cylon_table_t **out);
/* Unique / DistributedUnique (table.hpp:433,436). */
int cylon_unique(cylon_table_t *t, const int32_t *cols, int32_t ncols, int32_t distributed, cylon_table_t **out);
/* Shuffle (table.hpp:339). */
int cylon_shuffle(cylon_table_t *t, const int32_t *hash_cols, int32_t ncols, cylon_table_t **out);
/* Sort / DistributedSort (table.hpp:353-404): sort_cols[0] is the most significant column, ascending[c] != 0 = ascending.
 * num_bins / num_samples = SortOptions (0 = defaults: 16 bins per rank, 1 % of the rank's rows). */
int cylon_sort(cylon_table_t *t, const int32_t *sort_cols, const int32_t *ascending, int32_t ncols, int32_t distributed, uint32_t num_bins,
               uint64_t num_samples, cylon_table_t **out);
/* MapToSortPartitions (partition/partition.hpp:60-74): target partition per row and rows per partition into HOST arrays of
 * rows / num_partitions entries. Collective on a distributed context. */
int cylon_map_to_sort_partitions(cylon_table_t *t, int32_t column, uint32_t num_partitions, int32_t ascending, uint64_t num_samples, uint32_t num_bins,
                                 uint32_t *target_partitions /* host out */, uint32_t *partition_hist /* host out */);
/* HashPartition (table.hpp:348): out must hold num_partitions handles. */
int cylon_hash_partition(cylon_table_t *t, const int32_t *hash_cols, int32_t ncols, int32_t num_partitions, cylon_table_t **out);
/* HashGroupBy / DistributedHashGroupBy (groupby/groupby.hpp:23-33). ops: cyl_agg_op_t per aggregate column. */
int cylon_groupby(cylon_table_t *t, const int32_t *index_cols, int32_t nindex, const int32_t *agg_cols, const int32_t *agg_ops,
                  int32_t naggs, int32_t distributed, cylon_table_t **out);

#ifdef __cplusplus
}
#endif
#endif /* CYLON_B200_API_H_ */
