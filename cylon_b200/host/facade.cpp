// facade.cpp -- extern "C" handles over the C++ host API (include/cylon_b200_api.h); the counterpart of the
// reference's string-id facade cpp/src/cylon/table_api.cpp:117+ that its JNI and Cython bindings call.
#include "cylon_b200_api.h"

#include <algorithm>
#include <cstdarg>
#include <cstring>

#include "cylon/groupby/groupby.hpp"
#include "cylon/partition/partition.hpp"
#include "cylon/table.hpp"
#include "util.hpp"

namespace cylb {
int fail(int code, const char *fmt, ...);
}

struct cylon_context {
  std::shared_ptr<cylon::CylonContext> ctx;
};
struct cylon_table {
  std::shared_ptr<cylon::Table> table;
};

using cylon::Status;

static int report(const Status &st) {
  if (st.is_ok()) return CYL_OK;
  return cylb::fail(st.get_code(), "%s", st.get_msg().c_str());
}

static cylon_table *wrap(const std::shared_ptr<cylon::Table> &t) {
  auto *h = new cylon_table();
  h->table = t;
  return h;
}

// The C ids 0..3 and the reference's enum (SUM, MIN, MAX, COUNT, MEAN, VAR, NUNIQUE, QUANTILE, STDDEV) number alike, so an id
// passes through unchanged and the operator's own NotImplemented check answers for the ones outside the GPU path;
// anything that is not an enumerator at all is refused here.
static bool to_op(int32_t op, cylon::compute::AggregationOpId *out) {
  if (op < (int32_t)cylon::compute::SUM || op > (int32_t)cylon::compute::STDDEV) return false;
  *out = static_cast<cylon::compute::AggregationOpId>(op);
  return true;
}

extern "C" {

int cylon_context_init(int32_t device, cylon_context_t **out) {
  auto ctx = cylon::CylonContext::Init(device);
  if (!ctx) return cylb::fail(CYL_ExecutionError, "%s", cylon::CylonContext::LastError().c_str());
  *out = new cylon_context{ctx};
  return CYL_OK;
}

int cylon_nccl_unique_id(char id[128]) {
  cylon::net::NcclUniqueId u;
  Status st = cylon::net::NcclConfig::CreateUniqueId(&u);
  if (!st.is_ok()) return report(st);
  memcpy(id, u.bytes, 128);
  return CYL_OK;
}

int cylon_context_init_distributed(int32_t rank, int32_t world_size, int32_t device, const char id[128], cylon_context_t **out) {
  cylon::net::NcclUniqueId u;
  memcpy(u.bytes, id, 128);
  auto ctx = cylon::CylonContext::InitDistributed(cylon::net::NcclConfig::Make(rank, world_size, device, u));
  if (!ctx) return cylb::fail(CYL_ExecutionError, "%s", cylon::CylonContext::LastError().c_str());
  *out = new cylon_context{ctx};
  return CYL_OK;
}

int cylon_context_init_threads(const char *group, int32_t rank, int32_t world_size, int32_t device, cylon_context_t **out) {
  auto ctx = cylon::CylonContext::InitDistributed(cylon::net::ThreadConfig::Make(group ? group : "", rank, world_size, device));
  if (!ctx) return cylb::fail(CYL_ExecutionError, "%s", cylon::CylonContext::LastError().c_str());
  *out = new cylon_context{ctx};
  return CYL_OK;
}

int cylon_context_set_stream(cylon_context_t *c, void *stream) {
  c->ctx->SetStream(stream);
  return CYL_OK;
}
int cylon_context_add_config(cylon_context_t *c, const char *key, const char *value) {
  if (!c || !key || !value) return cylb::fail(CYL_Invalid, "cylon_context_add_config: null argument");
  c->ctx->AddConfig(key, value);
  return CYL_OK;
}
int cylon_context_rank(const cylon_context_t *c) { return c->ctx->GetRank(); }
int cylon_context_world_size(const cylon_context_t *c) { return c->ctx->GetWorldSize(); }
int cylon_context_barrier(cylon_context_t *c) {
  c->ctx->Barrier();
  return CYL_OK;
}
int cylon_context_finalize(cylon_context_t *c) {
  if (c) {
    c->ctx->Finalize();
    delete c;
  }
  return CYL_OK;
}

int cylon_table_from_host(cylon_context_t *c, const cyl_column_t *cols, const char *const *names, int32_t ncols, cylon_table_t **out) {
  std::vector<cyl_column_t> hc(cols, cols + ncols);
  std::vector<std::string> nm;
  for (int i = 0; i < ncols; i++) nm.push_back(names && names[i] ? names[i] : std::to_string(i));
  std::shared_ptr<cylon::Table> t;
  Status st = cylon::Table::FromHostColumns(c->ctx, hc, nm, t);
  if (!st.is_ok()) return report(st);
  *out = wrap(t);
  return CYL_OK;
}

int cylon_table_from_device(cylon_context_t *c, const cyl_column_t *cols, const char *const *names, int32_t ncols, cylon_table_t **out) {
  std::vector<std::shared_ptr<cylon::Column>> cv;
  for (int i = 0; i < ncols; i++) cv.push_back(cylon::Column::View(names && names[i] ? names[i] : std::to_string(i), cols[i]));
  std::shared_ptr<cylon::Table> t;
  Status st = cylon::Table::FromColumns(c->ctx, cv, t);
  if (!st.is_ok()) return report(st);
  *out = wrap(t);
  return CYL_OK;
}

int64_t cylon_table_rows(const cylon_table_t *t) { return t->table->Rows(); }
int32_t cylon_table_columns(const cylon_table_t *t) { return t->table->Columns(); }
const char *cylon_table_column_name(const cylon_table_t *t, int32_t col) {
  if (col < 0 || col >= t->table->Columns()) return "";
  return t->table->GetColumn(col)->name().c_str();
}

int cylon_table_column(const cylon_table_t *t, int32_t col, cyl_column_t *desc, int64_t *data_bytes) {
  if (col < 0 || col >= t->table->Columns()) return cylb::fail(CYL_IndexError, "column %d out of range", col);
  *desc = t->table->GetColumn(col)->desc();
  if (data_bytes) *data_bytes = t->table->GetColumn(col)->data_bytes();
  return CYL_OK;
}

int cylon_table_column_to_host(const cylon_table_t *t, int32_t col, void *data, int32_t *offsets, uint8_t *validity) {
  if (col < 0 || col >= t->table->Columns()) return cylb::fail(CYL_IndexError, "column %d out of range", col);
  return report(t->table->GetColumn(col)->ToHost(data, offsets, validity, t->table->GetContext()->GetStream()));
}

int cylon_table_retain_memory(cylon_table_t *t, int32_t retain) {
  t->table->retainMemory(retain != 0);
  return CYL_OK;
}

int cylon_table_checksum(const cylon_table_t *t, uint64_t *checksum) { return report(cylon::TableChecksum(t->table, checksum)); }

void cylon_table_free(cylon_table_t *t) { delete t; }

int cylon_join(cylon_table_t *l, cylon_table_t *r, int32_t join_type, int32_t algorithm, const int32_t *lc, const int32_t *rc, int32_t nkeys,
               const char *lp, const char *rp, int32_t distributed, cylon_table_t **out) {
  if (!l || !r) return cylb::fail(CYL_KeyError, "Couldn't find the %s table", l ? "right" : "left");
  if (nkeys < 1) return cylb::fail(CYL_Invalid, "join needs at least one key column");
  try {
    cylon::join::config::JoinConfig cfg((cylon::join::config::JoinType)join_type, std::vector<int>(lc, lc + nkeys), std::vector<int>(rc, rc + nkeys),
                                        (cylon::join::config::JoinAlgorithm)algorithm, lp ? lp : "", rp ? rp : "");
    std::shared_ptr<cylon::Table> o;
    Status st = distributed ? cylon::DistributedJoin(l->table, r->table, cfg, o) : cylon::Join(l->table, r->table, cfg, o);
    if (!st.is_ok()) return report(st);
    *out = wrap(o);
    return CYL_OK;
  } catch (const char *msg) {
    return cylb::fail(CYL_Invalid, "%s", msg);
  }
}

int cylon_set_operation(cylon_table_t *l, cylon_table_t *r, int32_t op, int32_t distributed, cylon_table_t **out) {
  if (!l || !r) return cylb::fail(CYL_KeyError, "Couldn't find the table");
  std::shared_ptr<cylon::Table> o;
  Status st;
  switch (op) {
    case CYL_SETOP_UNION: st = distributed ? cylon::DistributedUnion(l->table, r->table, o) : cylon::Union(l->table, r->table, o); break;
    case CYL_SETOP_SUBTRACT: st = distributed ? cylon::DistributedSubtract(l->table, r->table, o) : cylon::Subtract(l->table, r->table, o); break;
    case CYL_SETOP_INTERSECT: st = distributed ? cylon::DistributedIntersect(l->table, r->table, o) : cylon::Intersect(l->table, r->table, o); break;
    default: return cylb::fail(CYL_Invalid, "unknown set operation %d", op);
  }
  if (!st.is_ok()) return report(st);
  *out = wrap(o);
  return CYL_OK;
}

int cylon_unique(cylon_table_t *t, const int32_t *cols, int32_t ncols, int32_t distributed, cylon_table_t **out) {
  if (!t) return cylb::fail(CYL_KeyError, "Couldn't find the table");
  std::vector<int> c(cols, cols + ncols);
  std::shared_ptr<cylon::Table> o;
  Status st = distributed ? cylon::DistributedUnique(t->table, c, o) : cylon::Unique(t->table, c, o, true);
  if (!st.is_ok()) return report(st);
  *out = wrap(o);
  return CYL_OK;
}

int cylon_shuffle(cylon_table_t *t, const int32_t *cols, int32_t ncols, cylon_table_t **out) {
  if (!t) return cylb::fail(CYL_KeyError, "Couldn't find the table");
  std::vector<int> c(cols, cols + ncols);
  std::shared_ptr<cylon::Table> o;
  Status st = cylon::Shuffle(t->table, c, o);
  if (!st.is_ok()) return report(st);
  *out = wrap(o);
  return CYL_OK;
}

int cylon_sort(cylon_table_t *t, const int32_t *sort_cols, const int32_t *ascending, int32_t ncols, int32_t distributed, uint32_t num_bins,
               uint64_t num_samples, cylon_table_t **out) {
  if (!t) return cylb::fail(CYL_KeyError, "Couldn't find the table");
  if (ncols < 1 || !sort_cols || !ascending) return cylb::fail(CYL_Invalid, "cylon_sort: needs at least one sort column and its direction");
  std::vector<bool> dirs;
  for (int32_t i = 0; i < ncols; i++) dirs.push_back(ascending[i] != 0);
  std::shared_ptr<cylon::Table> o;
  Status st = distributed ? cylon::DistributedSort(t->table, std::vector<int>(sort_cols, sort_cols + ncols), o, dirs, cylon::SortOptions{num_bins, num_samples})
                          : cylon::Sort(t->table, std::vector<int32_t>(sort_cols, sort_cols + ncols), o, dirs);
  if (!st.is_ok()) return report(st);
  *out = wrap(o);
  return CYL_OK;
}

int cylon_map_to_sort_partitions(cylon_table_t *t, int32_t column, uint32_t num_partitions, int32_t ascending, uint64_t num_samples, uint32_t num_bins,
                                 uint32_t *target_partitions, uint32_t *partition_hist) {
  if (!t) return cylb::fail(CYL_KeyError, "Couldn't find the table");
  std::vector<uint32_t> pid, hist;
  Status st = cylon::MapToSortPartitions(t->table, column, num_partitions, pid, hist, ascending != 0, num_samples, num_bins);
  if (!st.is_ok()) return report(st);
  if (target_partitions) std::copy(pid.begin(), pid.end(), target_partitions);
  if (partition_hist) std::copy(hist.begin(), hist.end(), partition_hist);
  return CYL_OK;
}

int cylon_hash_partition(cylon_table_t *t, const int32_t *cols, int32_t ncols, int32_t num_partitions, cylon_table_t **out) {
  if (!t) return cylb::fail(CYL_KeyError, "Couldn't find the table");
  std::vector<int> c(cols, cols + ncols);
  std::unordered_map<int, std::shared_ptr<cylon::Table>> parts;
  Status st = cylon::HashPartition(t->table, c, num_partitions, &parts);
  if (!st.is_ok()) return report(st);
  for (int p = 0; p < num_partitions; p++) out[p] = wrap(parts[p]);
  return CYL_OK;
}

int cylon_groupby(cylon_table_t *t, const int32_t *index_cols, int32_t nindex, const int32_t *agg_cols, const int32_t *agg_ops, int32_t naggs,
                  int32_t distributed, cylon_table_t **out) {
  if (!t) return cylb::fail(CYL_KeyError, "Couldn't find the table");
  std::vector<int32_t> idx(index_cols, index_cols + nindex), ac(agg_cols, agg_cols + naggs);
  std::vector<cylon::compute::AggregationOpId> ops;
  for (int i = 0; i < naggs; i++) {
    cylon::compute::AggregationOpId op;
    if (!to_op(agg_ops[i], &op)) return cylb::fail(CYL_Invalid, "cylon_groupby: unknown aggregation id %d", agg_ops[i]);
    ops.push_back(op);
  }
  std::shared_ptr<cylon::Table> o;
  Status st;
  if (distributed) {
    st = cylon::DistributedHashGroupBy(t->table, idx, ac, ops, o);
  } else {
    std::vector<std::pair<int32_t, cylon::compute::AggregationOpId>> aggs;
    for (int i = 0; i < naggs; i++) aggs.emplace_back(ac[i], ops[i]);
    st = cylon::HashGroupBy(t->table, idx, aggs, o);
  }
  if (!st.is_ok()) return report(st);
  *out = wrap(o);
  return CYL_OK;
}

}  // extern "C"
