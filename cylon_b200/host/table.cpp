// table.cpp -- cylon::Table and the local operators: Join, Union/Subtract/Intersect, Unique.
//
// Replaces (reference, relative to cpp/src/cylon/): table.cpp:428-475 (Join), join/join.cpp:80-100 (JoinTables),
// join/join_utils.cpp:126-181 (build_final_table), table.cpp:531-721 (set operations), :923-982 (Unique). The
// control flow is the reference's; every data-touching step is a CUDA kernel reached through the C-ABI.
#include "cylon/table.hpp"

#include <algorithm>

#include "table_ops.hpp"

namespace cylon {

namespace detail {

std::vector<cyl_column_t> Descs(const Table &t, const std::vector<int> &cols) {
  std::vector<cyl_column_t> d;
  d.reserve(cols.size());
  for (int c : cols) d.push_back(t.GetColumn(c)->desc());
  return d;
}

std::vector<cyl_column_t> AllDescs(const Table &t) {
  std::vector<cyl_column_t> d;
  for (const auto &c : t.GetColumns()) d.push_back(c->desc());
  return d;
}

Status CheckColumns(const Table &t, const std::vector<int> &cols, const char *what) {
  if (cols.empty()) return Status(Code::Invalid, std::string(what) + ": no columns given");
  for (int c : cols)
    if (c < 0 || c >= t.Columns()) return Status(Code::IndexError, std::string(what) + ": column index " + std::to_string(c) + " out of range");
  return Status::OK();
}

// keep_source_validity: true for the Filter-like uses (set operators, Unique: arrow::compute::Filter keeps nulls); false
// for the join, whose copy_array_by_indices (util/copy_arrray.cpp:39-47) appends a null only for index -1 and otherwise
// the value slot as it is -- a null of the input comes out as a valid row holding whatever the slot held.
Status GatherColumns(const Table &t, const int64_t *idx, int64_t m, bool may_be_negative, bool keep_source_validity, const std::string &prefix,
                     void *stream, std::vector<std::shared_ptr<Column>> *out) {
  for (const auto &src : t.GetColumns()) {
    const cyl_column_t &d = src->desc();
    const uint8_t *src_validity = keep_source_validity ? d.validity : nullptr;
    const bool with_validity = may_be_negative || src_validity != nullptr;
    std::shared_ptr<Column> col;
    if (src->is_var_width()) {
      DevBuf offs;
      RETURN_CYLON_STATUS_IF_FAILED(offs.alloc((size_t)(m + 1) * 4, stream));
      int64_t total = 0;
      CYL_RETURN_IF_FAILED(cyl_gather_varwidth_offsets(d.offsets, idx, m, offs.as<int32_t>(), &total, stream));
      RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(prefix + src->name(), d.type, 0, m, total, with_validity, stream, &col));
      CYL_RETURN_IF_FAILED(cyl_memcpy_d2d(col->mutable_offsets(), offs.p, (m + 1) * 4, stream));
      CYL_RETURN_IF_FAILED(cyl_gather_varwidth_data(d.offsets, static_cast<const uint8_t *>(d.data), idx, m, col->desc().offsets,
                                                    static_cast<uint8_t *>(col->mutable_data()), stream));
    } else {
      RETURN_CYLON_STATUS_IF_FAILED(
          Column::Allocate(prefix + src->name(), d.type, d.byte_width, m, m * (int64_t)d.byte_width, with_validity, stream, &col));
      CYL_RETURN_IF_FAILED(cyl_gather_fixed(d.data, d.byte_width, idx, m, col->mutable_data(), nullptr, stream));
    }
    if (with_validity) CYL_RETURN_IF_FAILED(cyl_gather_validity(src_validity, idx, m, col->mutable_validity(), stream));
    out->push_back(col);
  }
  return Status::OK();
}

Status FilterTable(const std::shared_ptr<Table> &t, const uint8_t *mask, std::shared_ptr<Table> *out) {
  void *stream = t->GetContext()->GetStream();
  const int64_t n = t->Rows();
  DevBuf idx;
  RETURN_CYLON_STATUS_IF_FAILED(idx.alloc((size_t)(n + 1) * 8, stream));
  int64_t count = 0;
  CYL_RETURN_IF_FAILED(cyl_mask_to_indices(mask, n, idx.as<int64_t>(), &count, stream));
  std::vector<std::shared_ptr<Column>> cols;
  RETURN_CYLON_STATUS_IF_FAILED(GatherColumns(*t, idx.as<int64_t>(), count, false, true, "", stream, &cols));
  *out = std::make_shared<Table>(std::move(cols), t->GetContext());
  return Status::OK();
}

Status VerifyTableSchema(const Table &a, const Table &b) {  // table.cpp:213-236
  if (a.Columns() != b.Columns())
    return Status(Code::Invalid, "The no of columns of two tables are not similar. Can't perform union.");
  for (int c = 0; c < a.Columns(); c++)
    if (a.GetColumn(c)->type() != b.GetColumn(c)->type()) return Status(Code::Invalid, "The fields of two tables are not similar. Can't perform union.");
  return Status::OK();
}

Status ConcatTables(const std::shared_ptr<Table> &a, const std::shared_ptr<Table> &b, std::shared_ptr<Table> *out) {
  RETURN_CYLON_STATUS_IF_FAILED(VerifyTableSchema(*a, *b));
  void *stream = a->GetContext()->GetStream();
  const int64_t na = a->Rows(), nb = b->Rows();
  std::vector<std::shared_ptr<Column>> cols;
  for (int c = 0; c < a->Columns(); c++) {
    const auto &ca = a->GetColumn(c), &cb = b->GetColumn(c);
    const bool with_validity = ca->has_validity() || cb->has_validity();
    std::shared_ptr<Column> col;
    if (ca->is_var_width()) {
      const int64_t ba = ca->data_bytes(), bb = cb->data_bytes();
      RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(ca->name(), ca->type(), 0, na + nb, ba + bb, with_validity, stream, &col));
      CYL_RETURN_IF_FAILED(cyl_concat_offsets(ca->desc().offsets, na, cb->desc().offsets, nb, col->mutable_offsets(), stream));
      CYL_RETURN_IF_FAILED(cyl_memcpy_d2d(col->mutable_data(), ca->desc().data, ba, stream));
      CYL_RETURN_IF_FAILED(cyl_memcpy_d2d(static_cast<char *>(col->mutable_data()) + ba, cb->desc().data, bb, stream));
    } else {
      const int64_t w = ca->byte_width();
      RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(ca->name(), ca->type(), (int32_t)w, na + nb, (na + nb) * w, with_validity, stream, &col));
      CYL_RETURN_IF_FAILED(cyl_memcpy_d2d(col->mutable_data(), ca->desc().data, na * w, stream));
      CYL_RETURN_IF_FAILED(cyl_memcpy_d2d(static_cast<char *>(col->mutable_data()) + na * w, cb->desc().data, nb * w, stream));
    }
    if (with_validity) CYL_RETURN_IF_FAILED(cyl_concat_validity(ca->desc().validity, na, cb->desc().validity, nb, col->mutable_validity(), stream));
    cols.push_back(col);
  }
  *out = std::make_shared<Table>(std::move(cols), a->GetContext());
  return Status::OK();
}

}  // namespace detail

// ---------------------------------------------------------------------------------------------- Table
Status Table::FromColumns(std::shared_ptr<CylonContext> &ctx, std::vector<std::shared_ptr<Column>> columns, std::shared_ptr<Table> &out) {
  for (const auto &c : columns)
    if (c->length() != columns[0]->length()) return Status(Code::Invalid, "columns have different lengths");
  out = std::make_shared<Table>(std::move(columns), ctx);
  return Status::OK();
}

Status Table::FromHostColumns(std::shared_ptr<CylonContext> &ctx, const std::vector<cyl_column_t> &host_columns,
                              const std::vector<std::string> &names, std::shared_ptr<Table> &out) {
  std::vector<std::shared_ptr<Column>> cols;
  for (size_t i = 0; i < host_columns.size(); i++) {
    std::shared_ptr<Column> c;
    RETURN_CYLON_STATUS_IF_FAILED(Column::FromHost(i < names.size() ? names[i] : std::to_string(i), host_columns[i], ctx->GetStream(), &c));
    cols.push_back(c);
  }
  return FromColumns(ctx, std::move(cols), out);
}

std::vector<std::string> Table::ColumnNames() const {
  std::vector<std::string> n;
  for (const auto &c : columns_) n.push_back(c->name());
  return n;
}

Status Table::Project(const std::vector<int32_t> &cols, std::shared_ptr<Table> &out) const {
  std::vector<std::shared_ptr<Column>> sel;
  for (int32_t c : cols) {
    if (c < 0 || c >= Columns()) return Status(Code::IndexError, "Project: column index out of range");
    sel.push_back(columns_[c]);
  }
  out = std::make_shared<Table>(std::move(sel), ctx_);
  return Status::OK();
}

// ---------------------------------------------------------------------------------------------- Join
Status Join(std::shared_ptr<Table> &left, std::shared_ptr<Table> &right, const join::config::JoinConfig &cfg, std::shared_ptr<Table> &out) {
  if (left == nullptr) return Status(Code::KeyError, "Couldn't find the left table");    // table.cpp:430-431
  if (right == nullptr) return Status(Code::KeyError, "Couldn't find the right table");  // table.cpp:432-433
  const auto &lidx = cfg.GetLeftColumnIdx(), &ridx = cfg.GetRightColumnIdx();
  if (lidx.size() != ridx.size()) return Status(Code::Invalid, "left and right column indices sizes are not equal");
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*left, lidx, "Join(left)"));
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*right, ridx, "Join(right)"));
  auto ctx = left->GetContext();
  void *stream = ctx->GetStream();
  const auto lk = detail::Descs(*left, lidx), rk = detail::Descs(*right, ridx);
  const int jt = (int)cfg.GetType();

  // fast path: radix-partitioned shared-memory join that writes the result columns directly. JoinAlgorithm does not
  // matter here: HASH and SORT are two CPU strategies for the same row multiset (join/join.cpp:80-100); the LSD radix
  // sort + merge of join_sort.cu remains the general SORT path ("cylon.join.fused" = "0", strings, wide tables).
  if (lidx.size() == 1 && ctx->GetConfig("cylon.join.fused", "1") != "0") {
    const auto lall = detail::AllDescs(*left), rall = detail::AllDescs(*right);
    if (cyl_hash_join_fused_supported(lall.data(), (int32_t)lall.size(), lidx[0], rall.data(), (int32_t)rall.size(), ridx[0], jt)) {
      const int64_t nl = left->Rows(), nr = right->Rows();
      int64_t capacity = std::max(nl, nr) + std::max(nl, nr) / 8 + 1024;
      for (int attempt = 0; attempt < 2; attempt++) {
        std::vector<std::shared_ptr<Column>> cols;
        std::vector<void *> outs;
        for (int side = 0; side < 2; side++) {
          const Table &t = side == 0 ? *left : *right;
          const std::string &prefix = side == 0 ? cfg.GetLeftTableSuffix() : cfg.GetRightTableSuffix();
          for (const auto &src : t.GetColumns()) {
            std::shared_ptr<Column> col;
            RETURN_CYLON_STATUS_IF_FAILED(
                Column::Allocate(prefix + src->name(), src->type(), src->byte_width(), capacity, capacity * (int64_t)src->byte_width(), false, stream, &col));
            outs.push_back(col->mutable_data());
            cols.push_back(col);
          }
        }
        int64_t m = 0;
        CYL_RETURN_IF_FAILED(cyl_hash_join_fused_inner(lall.data(), (int32_t)lall.size(), lidx[0], rall.data(), (int32_t)rall.size(), ridx[0], nl, nr,
                                                       capacity, outs.data(), &m, stream));
        if (m <= capacity) {
          for (auto &c : cols) c->truncate(m);
          out = std::make_shared<Table>(std::move(cols), ctx);
          return Status::OK();
        }
        capacity = m;  // the optimistic capacity was too small: the exact size is known now
      }
      return Status(Code::ExecutionError, "fused join did not fit its exact capacity");
    }
  }

  // second fast path (join_fused8x.cu): outer joins, two-column keys, 4/8-byte columns -- for HASH and SORT alike, since
  // both algorithms of the reference produce the same row multiset (join/join.cpp:80-100)
  if ((lidx.size() == 1 || lidx.size() == 2) && ctx->GetConfig("cylon.join.fused", "1") != "0") {
    const auto lall = detail::AllDescs(*left), rall = detail::AllDescs(*right);
    std::vector<int32_t> lk32(lidx.begin(), lidx.end()), rk32(ridx.begin(), ridx.end());
    if (cyl_join_fused_supported(lall.data(), (int32_t)lall.size(), lk32.data(), rall.data(), (int32_t)rall.size(), rk32.data(), (int32_t)lk32.size(), jt)) {
      const int64_t nl = left->Rows(), nr = right->Rows();
      const bool left_null = jt == CYL_JOIN_RIGHT || jt == CYL_JOIN_FULL_OUTER;
      const bool right_null = jt == CYL_JOIN_LEFT || jt == CYL_JOIN_FULL_OUTER;
      // optimistic result size: one partner per row for INNER; an outer join also keeps every unmatched row of its
      // preserved side(s), so all rows of both tables (exact when keys are unique on one side, the usual shape)
      int64_t capacity = std::max(nl, nr) + std::max(nl, nr) / 8 + 1024;
      if (jt != CYL_JOIN_INNER) capacity = nl + nr + 1024;
      for (int attempt = 0; attempt < 2; attempt++) {
        std::vector<std::shared_ptr<Column>> cols;
        std::vector<void *> outs;
        for (int side = 0; side < 2; side++) {
          const Table &t = side == 0 ? *left : *right;
          const std::string &prefix = side == 0 ? cfg.GetLeftTableSuffix() : cfg.GetRightTableSuffix();
          const bool nullable = side == 0 ? left_null : right_null;
          for (const auto &src : t.GetColumns()) {
            std::shared_ptr<Column> col;
            RETURN_CYLON_STATUS_IF_FAILED(Column::Allocate(prefix + src->name(), src->type(), src->byte_width(), capacity,
                                                           capacity * (int64_t)src->byte_width(), nullable, stream, &col));
            outs.push_back(col->mutable_data());
            cols.push_back(col);
          }
        }
        detail::DevBuf lvalid, rvalid;
        if (left_null) RETURN_CYLON_STATUS_IF_FAILED(lvalid.alloc((size_t)capacity + 16, stream));
        if (right_null) RETURN_CYLON_STATUS_IF_FAILED(rvalid.alloc((size_t)capacity + 16, stream));
        int64_t m = 0;
        CYL_RETURN_IF_FAILED(cyl_join_fused(lall.data(), (int32_t)lall.size(), lk32.data(), rall.data(), (int32_t)rall.size(), rk32.data(),
                                            (int32_t)lk32.size(), jt, capacity, outs.data(), left_null ? lvalid.as<uint8_t>() : nullptr,
                                            right_null ? rvalid.as<uint8_t>() : nullptr, &m, stream));
        if (m <= capacity) {
          for (size_t c = 0; c < cols.size(); c++) {
            cols[c]->truncate(m);
            const bool is_left = c < lall.size();
            if ((is_left && left_null) || (!is_left && right_null))
              CYL_RETURN_IF_FAILED(cyl_pack_validity(is_left ? lvalid.as<uint8_t>() : rvalid.as<uint8_t>(), m, cols[c]->mutable_validity(), stream));
          }
          out = std::make_shared<Table>(std::move(cols), ctx);
          return Status::OK();
        }
        capacity = m;
      }
      return Status(Code::ExecutionError, "fused join did not fit its exact capacity");
    }
  }

  cyl_join_plan_t *plan = nullptr;
  int64_t m = 0;
  if (cfg.GetAlgorithm() == join::config::HASH) {  // join/join.cpp:86-92
    CYL_RETURN_IF_FAILED(cyl_hash_join_count(lk.data(), rk.data(), (int32_t)lk.size(), left->Rows(), right->Rows(), jt, &plan, &m, stream));
  } else {
    CYL_RETURN_IF_FAILED(cyl_sort_join_count(lk.data(), rk.data(), (int32_t)lk.size(), left->Rows(), right->Rows(), jt, &plan, &m, stream));
  }
  detail::DevBuf li, ri;
  Status st = li.alloc((size_t)(m + 1) * 8, stream);
  if (st.is_ok()) st = ri.alloc((size_t)(m + 1) * 8, stream);
  if (!st.is_ok()) {
    cyl_join_plan_free(plan);
    return st;
  }
  CYL_RETURN_IF_FAILED(cyl_join_materialize(plan, li.as<int64_t>(), ri.as<int64_t>(), stream));

  // build_final_table (join_utils.cpp:126-181): all left columns then all right columns, names prefix + name
  const bool left_null = jt == CYL_JOIN_RIGHT || jt == CYL_JOIN_FULL_OUTER;
  const bool right_null = jt == CYL_JOIN_LEFT || jt == CYL_JOIN_FULL_OUTER;
  std::vector<std::shared_ptr<Column>> cols;
  RETURN_CYLON_STATUS_IF_FAILED(detail::GatherColumns(*left, li.as<int64_t>(), m, left_null, false, cfg.GetLeftTableSuffix(), stream, &cols));
  RETURN_CYLON_STATUS_IF_FAILED(detail::GatherColumns(*right, ri.as<int64_t>(), m, right_null, false, cfg.GetRightTableSuffix(), stream, &cols));
  out = std::make_shared<Table>(std::move(cols), ctx);
  return Status::OK();
}

// ---------------------------------------------------------------------------------------------- set operations
static Status LocalSetOp(std::shared_ptr<Table> &first, std::shared_ptr<Table> &second, int op, std::shared_ptr<Table> &out) {
  if (first == nullptr || second == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  RETURN_CYLON_STATUS_IF_FAILED(detail::VerifyTableSchema(*first, *second));
  auto ctx = first->GetContext();
  void *stream = ctx->GetStream();
  const int64_t nl = first->Rows(), nr = second->Rows();
  detail::DevBuf ml, mr;
  RETURN_CYLON_STATUS_IF_FAILED(ml.alloc((size_t)nl + 1, stream));
  RETURN_CYLON_STATUS_IF_FAILED(mr.alloc((size_t)nr + 1, stream));
  const auto lc = detail::AllDescs(*first), rc = detail::AllDescs(*second);
  CYL_RETURN_IF_FAILED(cyl_rowset_mask(lc.data(), rc.data(), (int32_t)lc.size(), nl, nr, op, ml.as<uint8_t>(), mr.as<uint8_t>(), stream));
  std::shared_ptr<Table> fl;
  RETURN_CYLON_STATUS_IF_FAILED(detail::FilterTable(first, ml.as<uint8_t>(), &fl));
  if (op != CYL_SETOP_UNION) {
    out = fl;
    return Status::OK();
  }
  std::shared_ptr<Table> fr;
  RETURN_CYLON_STATUS_IF_FAILED(detail::FilterTable(second, mr.as<uint8_t>(), &fr));
  return detail::ConcatTables(fl, fr, &out);  // table.cpp:590-600
}

Status Union(std::shared_ptr<Table> &first, std::shared_ptr<Table> &second, std::shared_ptr<Table> &out) {
  return LocalSetOp(first, second, CYL_SETOP_UNION, out);
}
Status Subtract(std::shared_ptr<Table> &first, std::shared_ptr<Table> &second, std::shared_ptr<Table> &out) {
  return LocalSetOp(first, second, CYL_SETOP_SUBTRACT, out);
}
Status Intersect(std::shared_ptr<Table> &first, std::shared_ptr<Table> &second, std::shared_ptr<Table> &out) {
  return LocalSetOp(first, second, CYL_SETOP_INTERSECT, out);
}

Status Unique(std::shared_ptr<Table> &in, const std::vector<int> &cols, std::shared_ptr<Table> &out, bool first) {
  if (in == nullptr) return Status(Code::KeyError, "Couldn't find the table");
  if (!first) return Status(Code::NotImplemented, "Unique(first=false) is not supported (the reference's branch is broken, table.cpp:953-958)");
  RETURN_CYLON_STATUS_IF_FAILED(detail::CheckColumns(*in, cols, "Unique"));
  void *stream = in->GetContext()->GetStream();
  const int64_t n = in->Rows();
  detail::DevBuf mask;
  RETURN_CYLON_STATUS_IF_FAILED(mask.alloc((size_t)n + 1, stream));
  const auto kc = detail::Descs(*in, cols);
  CYL_RETURN_IF_FAILED(cyl_rowset_mask(kc.data(), nullptr, (int32_t)kc.size(), n, 0, CYL_SETOP_UNIQUE, mask.as<uint8_t>(), nullptr, stream));
  return detail::FilterTable(in, mask.as<uint8_t>(), &out);
}

Status TableChecksum(const std::shared_ptr<Table> &table, uint64_t *checksum) {
  const auto d = detail::AllDescs(*table);
  CYL_RETURN_IF_FAILED(cyl_table_checksum(d.data(), (int32_t)d.size(), table->Rows(), checksum, table->GetContext()->GetStream()));
  return Status::OK();
}

}  // namespace cylon
