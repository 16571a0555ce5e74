// groupby.cpp -- HashGroupBy / DistributedHashGroupBy on the device.
//
// Replaces (reference, relative to cpp/src/cylon/): groupby/hash_groupby.cpp:250-360 and groupby/groupby.cpp:24-91.
// Same control flow: project (index ++ aggregate columns), pre-aggregate locally when world == 1 or every op is in
// {SUM, MIN, MAX} (groupby.cpp:24-31,60-67), shuffle on the index columns, aggregate again -- which is also why the
// output column of a distributed associative aggregation is named "sum_sum_x" (SURVEY.md G1); kept for parity.
#include "cylon/groupby/groupby.hpp"

#include <algorithm>

#include "table_ops.hpp"

namespace cylon {

Status HashGroupBy(const std::shared_ptr<Table> &table, const std::vector<int32_t> &idx_cols,
                   const std::vector<std::pair<int32_t, compute::AggregationOpId>> &aggregations, std::shared_ptr<Table> &output) {
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  std::vector<int> keys(idx_cols.begin(), idx_cols.end());
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*table, keys, "HashGroupBy"));
  for (const auto &a : aggregations) {
    if (a.first < 0 || a.first >= table->Columns()) return Status(Code::IndexError, "HashGroupBy: aggregate column out of range");
    if (a.second == compute::NUNIQUE || a.second == compute::QUANTILE || (int)a.second < 0 || (int)a.second > (int)compute::STDDEV)
      return Status(Code::NotImplemented, "HashGroupBy: NUNIQUE / QUANTILE do not run on the GPU path");
  }
  auto ctx = table->GetContext();
  void *stream = ctx->GetStream();
  const int64_t n = table->Rows();

  // fast path: one 8-byte key, 8-byte numeric values -> groups and all aggregates in one sweep (groupby_fused8.cu)
  if (idx_cols.size() == 1 && ctx->GetConfig("cylon.groupby.fused", "1") != "0") {
    const cyl_column_t key = table->GetColumn(idx_cols[0])->desc();
    std::vector<cyl_column_t> values;  // distinct source columns (the same column may feed several aggregates)
    std::vector<int32_t> agg_value, agg_op;
    for (const auto &a : aggregations) {
      const cyl_column_t d = table->GetColumn(a.first)->desc();
      int32_t at = -1;
      for (size_t v = 0; v < values.size(); v++)
        if (values[v].data == d.data && values[v].type == d.type) at = (int32_t)v;
      if (at < 0) {
        values.push_back(d);
        at = (int32_t)values.size() - 1;
      }
      agg_value.push_back(at);
      agg_op.push_back((int32_t)a.second);  // AggregationOpId and cyl_agg_op_t number alike; the fused path takes the first four
    }
    if (n > 0 && cyl_groupby_fused_supported(&key, values.data(), (int32_t)values.size(), agg_value.data(), agg_op.data(), (int32_t)agg_op.size())) {
      int64_t capacity = std::min<int64_t>(n, n / 4 + (1 << 20));
      bool fused_failed = false;
      for (int attempt = 0; attempt < 2; attempt++) {
        std::vector<std::shared_ptr<Column>> cols;
        std::shared_ptr<Column> kc;
        const auto &ksrc = table->GetColumn(idx_cols[0]);
        RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(ksrc->name(), ksrc->type(), 8, capacity, capacity * 8, false, stream, &kc));
        cols.push_back(kc);
        std::vector<void *> outs;
        for (const auto &a : aggregations) {
          const auto &src = table->GetColumn(a.first);
          const int32_t type = a.second == compute::COUNT ? (int32_t)CYL_INT64 : src->type();
          std::shared_ptr<Column> col;
          RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(std::string(compute::AggregationPrefix(a.second)) + src->name(), type, 8, capacity, capacity * 8,
                                                         false, stream, &col));
          outs.push_back(col->mutable_data());
          cols.push_back(col);
        }
        int64_t groups = 0;
        const int rc = cyl_groupby_fused(&key, values.data(), (int32_t)values.size(), agg_value.data(), agg_op.data(), (int32_t)agg_op.size(),
                                         capacity, kc->mutable_data(), outs.data(), &groups, stream);
        if (rc == CYL_CapacityError) {  // keys that pile into one bin of the shared-memory table: the general path has no such limit
          fused_failed = true;
          break;
        }
        if (rc != CYL_OK) return detail::FromCode(rc);
        if (groups <= capacity) {
          for (auto &c : cols) c->truncate(groups);
          output = std::make_shared<Table>(std::move(cols), ctx);
          return Status::OK();
        }
        capacity = groups;  // the optimistic capacity was too small: the exact count is known now
      }
      if (!fused_failed) return Status(Code::ExecutionError, "fused group-by did not fit its exact capacity");
    }
  }

  // make_groups (hash_groupby.cpp:92-126)
  detail::DevBuf gid, first;
  RETURN_CYLON_STATUS_IF_FAILED(gid.alloc((size_t)(n + 1) * 8, stream));
  RETURN_CYLON_STATUS_IF_FAILED(first.alloc((size_t)n + 1, stream));
  const auto kd = detail::Descs(*table, keys);
  int64_t groups = 0;
  CYL_RETURN_IF_FAILED(cyl_groupby_hash(kd.data(), (int32_t)kd.size(), n, gid.as<int64_t>(), first.as<uint8_t>(), &groups, stream));

  // key columns = first row of every group (hash_groupby.cpp:281-286)
  std::shared_ptr<Table> key_table, filtered;
  RETURN_CYLON_STATUS_IF_FAILED(table->Project(idx_cols, key_table));
  RETURN_CYLON_STATUS_IF_FAILED(detail::FilterTable(key_table, first.as<uint8_t>(), &filtered));
  std::vector<std::shared_ptr<Column>> cols = filtered->GetColumns();

  // aggregates (hash_groupby.cpp:288-306); output name = prefix + input name (:197-199)
  for (const auto &a : aggregations) {
    const auto &src = table->GetColumn(a.first);
    const bool count = a.second == compute::COUNT;
    const bool as_double = a.second == compute::VAR || a.second == compute::STDDEV;  // KernelTraits<VAR,T>::ResultT = double
    const int32_t type = count ? (int32_t)CYL_INT64 : (as_double ? (int32_t)CYL_DOUBLE : src->type());
    const int32_t width = (count || as_double) ? 8 : src->byte_width();
    std::shared_ptr<Column> col;
    RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(std::string(compute::AggregationPrefix(a.second)) + src->name(), type, width, groups,
                                                   groups * (int64_t)width, false, stream, &col));
    const int op = (int)a.second;
    CYL_RETURN_IF_FAILED(cyl_aggregate(&src->desc(), gid.as<int64_t>(), groups, op, col->mutable_data(), stream));
    cols.push_back(col);
  }
  output = std::make_shared<Table>(std::move(cols), ctx);
  return Status::OK();
}

Status HashGroupBy(const std::shared_ptr<Table> &table, int32_t idx_col, const std::vector<int32_t> &aggregate_cols,
                   const std::vector<compute::AggregationOpId> &aggregate_ops, std::shared_ptr<Table> &output) {
  if (aggregate_cols.size() != aggregate_ops.size()) return Status(Code::Invalid, "aggregate_cols size != aggregate_ops size");
  std::vector<std::pair<int32_t, compute::AggregationOpId>> aggs;
  for (size_t i = 0; i < aggregate_cols.size(); i++) aggs.emplace_back(aggregate_cols[i], aggregate_ops[i]);
  return HashGroupBy(table, std::vector<int32_t>{idx_col}, aggs, output);
}

static bool is_associative(const std::vector<compute::AggregationOpId> &ops) {  // groupby.cpp:24-31
  return std::all_of(ops.begin(), ops.end(), [](compute::AggregationOpId op) { return op == compute::SUM || op == compute::MIN || op == compute::MAX; });
}

Status DistributedHashGroupBy(std::shared_ptr<Table> &table, const std::vector<int32_t> &index_cols, const std::vector<int32_t> &aggregate_cols,
                              const std::vector<compute::AggregationOpId> &aggregate_ops, std::shared_ptr<Table> &output) {
  if (table == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  if (aggregate_cols.size() != aggregate_ops.size()) return Status(Code::Invalid, "aggregate_cols size != aggregate_ops size");  // groupby.cpp:38-40
  std::vector<int32_t> project_cols(index_cols);
  project_cols.insert(project_cols.end(), aggregate_cols.begin(), aggregate_cols.end());
  std::shared_ptr<Table> projected;
  RETURN_CYLON_STATUS_IF_FAILED(table->Project(project_cols, projected));  // groupby.cpp:43-46
  std::vector<int32_t> idx_after(index_cols.size());
  for (size_t i = 0; i < idx_after.size(); i++) idx_after[i] = (int32_t)i;
  std::vector<std::pair<int32_t, compute::AggregationOpId>> aggs;
  for (size_t i = 0; i < aggregate_cols.size(); i++) aggs.emplace_back((int32_t)(index_cols.size() + i), aggregate_ops[i]);

  const int world = table->GetContext()->GetWorldSize();
  bool basic_ops = true;  // SUM / MIN / MAX / COUNT only: partial results can be combined after the shuffle
  for (auto op : aggregate_ops) {  // refuse before anything is shuffled (the reference finds out after its shuffle, hash_groupby.cpp:217-248)
    if (op == compute::NUNIQUE || op == compute::QUANTILE || (int)op < 0 || (int)op > (int)compute::STDDEV)
      return Status(Code::NotImplemented, "DistributedHashGroupBy: NUNIQUE / QUANTILE do not run on the GPU path");
    basic_ops = basic_ops && (op == compute::SUM || op == compute::MIN || op == compute::MAX || op == compute::COUNT);
  }
  std::shared_ptr<Table> local;
  if (world > 1 && basic_ops && !is_associative(aggregate_ops) && table->GetContext()->GetConfig("cylon.groupby.preaggregate", "1") != "0") {
    // COUNT among the operations. The reference then ships every raw row through the shuffle (groupby.cpp:60-77: it only
    // pre-aggregates when all operations are in {SUM, MIN, MAX}). Here the rows are still reduced locally first and the
    // partial COUNTs are SUMmed after the shuffle, so only one row per (rank, group) crosses NVLink; the result columns
    // get the names the reference's single application produces (count_x, not sum_count_x). Values are identical except
    // for the association order of floating-point SUMs (two levels instead of one; within the 1e-12 tolerance).
    RETURN_CYLON_STATUS_IF_FAILED(HashGroupBy(projected, idx_after, aggs, local));
    std::shared_ptr<Table> shuffled;
    std::vector<int> hc(idx_after.begin(), idx_after.end());
    RETURN_CYLON_STATUS_IF_FAILED(Shuffle(local, hc, shuffled));
    std::vector<std::pair<int32_t, compute::AggregationOpId>> combine;
    for (size_t i = 0; i < aggregate_ops.size(); i++)
      combine.emplace_back((int32_t)(index_cols.size() + i), aggregate_ops[i] == compute::COUNT ? compute::SUM : aggregate_ops[i]);
    RETURN_CYLON_STATUS_IF_FAILED(HashGroupBy(shuffled, idx_after, combine, output));
    for (size_t i = 0; i < aggregate_ops.size(); i++)
      output->GetColumn((int)(index_cols.size() + i))
          ->set_name(std::string(compute::AggregationPrefix(aggregate_ops[i])) + projected->GetColumn((int)(index_cols.size() + i))->name());
    return Status::OK();
  }
  if (world == 1 || is_associative(aggregate_ops)) {  // groupby.cpp:60-64
    RETURN_CYLON_STATUS_IF_FAILED(HashGroupBy(projected, idx_after, aggs, local));
  } else {
    local = projected;
  }
  if (world > 1) {  // groupby.cpp:69-77
    std::shared_ptr<Table> shuffled;
    std::vector<int> hc(idx_after.begin(), idx_after.end());
    RETURN_CYLON_STATUS_IF_FAILED(Shuffle(local, hc, shuffled));
    return HashGroupBy(shuffled, idx_after, aggs, output);
  }
  output = local;
  return Status::OK();
}

Status DistributedHashGroupBy(std::shared_ptr<Table> &table, int32_t index_col, const std::vector<int32_t> &aggregate_cols,
                              const std::vector<compute::AggregationOpId> &aggregate_ops, std::shared_ptr<Table> &output) {
  return DistributedHashGroupBy(table, std::vector<int32_t>{index_col}, aggregate_cols, aggregate_ops, output);
}

}  // namespace cylon
