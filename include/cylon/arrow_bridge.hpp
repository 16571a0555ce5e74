// Arrow host tables in and out of HBM: Table::FromArrowTable / ToArrowTable / get_table with the reference's signatures
// (cpp/src/cylon/table.hpp:76,96,146; table.cpp:202-211,477-480). Header-only and optional: include it from code that is
// compiled against Arrow C++ (any recent version; link libarrow). libcylon_b200.so itself does not depend on Arrow.
//
//   #include <cylon/arrow_bridge.hpp>
//   std::shared_ptr<cylon::Table> t;
//   cylon::Table::FromArrowTable(ctx, arrow_table, t);     // validity / offsets / data buffers -> HBM, as they are
//   ... cylon::DistributedJoin(...), cylon::HashGroupBy(...) ...
//   std::shared_ptr<arrow::Table> host;  out->ToArrowTable(host);
//
// Supported column types: the integer, floating-point, date, time and timestamp types, string, binary and
// fixed_size_binary -- the types the reference's partition / split / comparator kernels switch over
// (arrow/arrow_partition_kernels.cpp:307-330). Anything else returns Code::NotImplemented. Chunked columns are
// combined first (the reference's operators require single-chunk tables too, table.cpp:535,610,669).
#pragma once
#include <arrow/api.h>
#include <arrow/util/bitmap_ops.h>

#include <memory>
#include <string>
#include <vector>

#include "../cylon_b200.h"
#include "table.hpp"

namespace cylon {
namespace arrow_bridge {

inline bool FixedWidthOf(const std::shared_ptr<arrow::DataType> &t, int32_t *width) {
  switch (t->id()) {
    case arrow::Type::UINT8: case arrow::Type::INT8: *width = 1; return true;
    case arrow::Type::UINT16: case arrow::Type::INT16: *width = 2; return true;
    case arrow::Type::UINT32: case arrow::Type::INT32: case arrow::Type::FLOAT: case arrow::Type::DATE32: case arrow::Type::TIME32:
      *width = 4; return true;
    case arrow::Type::UINT64: case arrow::Type::INT64: case arrow::Type::DOUBLE: case arrow::Type::DATE64: case arrow::Type::TIMESTAMP:
    case arrow::Type::TIME64: *width = 8; return true;
    case arrow::Type::FIXED_SIZE_BINARY: *width = static_cast<const arrow::FixedSizeBinaryType &>(*t).byte_width(); return true;
    default: return false;
  }
}

inline std::shared_ptr<arrow::DataType> TypeOf(int32_t id, int32_t byte_width) {
  switch (id) {
    case CYL_UINT8: return arrow::uint8();
    case CYL_INT8: return arrow::int8();
    case CYL_UINT16: return arrow::uint16();
    case CYL_INT16: return arrow::int16();
    case CYL_UINT32: return arrow::uint32();
    case CYL_INT32: return arrow::int32();
    case CYL_UINT64: return arrow::uint64();
    case CYL_INT64: return arrow::int64();
    case CYL_FLOAT: return arrow::float32();
    case CYL_DOUBLE: return arrow::float64();
    case CYL_STRING: return arrow::utf8();
    case CYL_BINARY: return arrow::binary();
    case CYL_FIXED_SIZE_BINARY: return arrow::fixed_size_binary(byte_width);
    case CYL_DATE32: return arrow::date32();
    case CYL_DATE64: return arrow::date64();
    case CYL_TIMESTAMP: return arrow::timestamp(arrow::TimeUnit::MICRO);  // the unit is not carried through the device table
    case CYL_TIME32: return arrow::time32(arrow::TimeUnit::MILLI);
    case CYL_TIME64: return arrow::time64(arrow::TimeUnit::MICRO);
    default: return nullptr;
  }
}

}  // namespace arrow_bridge

inline Status Table::FromArrowTable(std::shared_ptr<CylonContext> &ctx, std::shared_ptr<arrow::Table> &table, std::shared_ptr<Table> &out) {
  if (!table) return Status(Code::Invalid, "FromArrowTable: null table");
  auto combined = table->CombineChunks();
  if (!combined.ok()) return Status(Code::ExecutionError, combined.status().ToString());
  const std::shared_ptr<arrow::Table> &t = *combined;
  std::vector<cyl_column_t> host((size_t)t->num_columns());
  std::vector<std::string> names;
  std::vector<std::shared_ptr<arrow::Buffer>> keep;  // re-aligned bitmaps / offsets live until the copies are done
  for (int c = 0; c < t->num_columns(); c++) {
    names.push_back(t->field(c)->name());
    cyl_column_t &h = host[(size_t)c];
    h = cyl_column_t{};
    const auto &chunked = t->column(c);
    const auto type = chunked->type();
    h.type = (int32_t)type->id();
    if (chunked->num_chunks() == 0 || chunked->length() == 0) {  // empty column: a descriptor without buffers
      int32_t w = 0;
      const bool var = type->id() == arrow::Type::STRING || type->id() == arrow::Type::BINARY;
      if (!var && !arrow_bridge::FixedWidthOf(type, &w)) return Status(Code::NotImplemented, "FromArrowTable: unsupported type " + type->ToString());
      static const int32_t zero_offset[1] = {0};
      h.byte_width = w;
      h.length = 0;
      h.offsets = var ? zero_offset : nullptr;
      continue;
    }
    const std::shared_ptr<arrow::ArrayData> &d = chunked->chunk(0)->data();
    const int64_t off = d->offset, n = d->length;
    h.length = n;
    h.null_count = d->GetNullCount();
    if (h.null_count > 0 && d->buffers[0]) {
      if (off % 8 == 0) {
        h.validity = d->buffers[0]->data() + off / 8;
      } else {  // a sliced array: bring bit `off` to bit 0
        auto copy = arrow::internal::CopyBitmap(arrow::default_memory_pool(), d->buffers[0]->data(), off, n);
        if (!copy.ok()) return Status(Code::ExecutionError, copy.status().ToString());
        keep.push_back(*copy);
        h.validity = keep.back()->data();
      }
    }
    if (type->id() == arrow::Type::STRING || type->id() == arrow::Type::BINARY) {
      h.byte_width = 0;
      h.offsets = reinterpret_cast<const int32_t *>(d->buffers[1]->data()) + off;  // Column::FromHost rebases offsets[0] != 0
      h.data = d->buffers[2] ? d->buffers[2]->data() : nullptr;
    } else {
      int32_t w = 0;
      if (!arrow_bridge::FixedWidthOf(type, &w)) return Status(Code::NotImplemented, "FromArrowTable: unsupported type " + type->ToString());
      h.byte_width = w;
      h.data = d->buffers[1]->data() + off * (int64_t)w;
    }
  }
  RETURN_CYLON_STATUS_IF_FAILED(FromHostColumns(ctx, host, names, out));
  if (cyl_stream_sync(ctx->GetStream()) != CYL_OK) return Status(Code::ExecutionError, cyl_last_error());  // Arrow buffers may go away now
  return Status::OK();
}

inline Status Table::ToArrowTable(std::shared_ptr<arrow::Table> &output) {
  std::vector<std::shared_ptr<arrow::Field>> fields;
  std::vector<std::shared_ptr<arrow::Array>> arrays;
  void *stream = ctx_->GetStream();
  for (const auto &col : columns_) {
    const cyl_column_t &d = col->desc();
    auto type = arrow_bridge::TypeOf(d.type, d.byte_width);
    if (!type) return Status(Code::NotImplemented, "ToArrowTable: column type id " + std::to_string(d.type) + " has no Arrow mapping here");
    const int64_t n = d.length;
    std::shared_ptr<arrow::Buffer> validity, offsets, data;
    auto alloc = [](int64_t bytes, std::shared_ptr<arrow::Buffer> *out_buf) -> Status {
      auto r = arrow::AllocateBuffer(bytes < 8 ? 8 : bytes);
      if (!r.ok()) return Status(Code::OutOfMemory, r.status().ToString());
      *out_buf = std::move(*r);
      return Status::OK();
    };
    RETURN_CYLON_STATUS_IF_FAILED(alloc(col->data_bytes(), &data));
    if (col->is_var_width()) RETURN_CYLON_STATUS_IF_FAILED(alloc((n + 1) * 4, &offsets));
    if (d.validity) RETURN_CYLON_STATUS_IF_FAILED(alloc((n + 7) / 8 + 8, &validity));
    if (col->is_var_width() && n == 0) reinterpret_cast<int32_t *>(offsets->mutable_data())[0] = 0;
    RETURN_CYLON_STATUS_IF_FAILED(col->ToHost(data->mutable_data(), offsets ? reinterpret_cast<int32_t *>(offsets->mutable_data()) : nullptr,
                                              validity ? validity->mutable_data() : nullptr, stream));
    std::vector<std::shared_ptr<arrow::Buffer>> bufs;
    bufs.push_back(validity);
    if (col->is_var_width()) bufs.push_back(offsets);
    bufs.push_back(data);
    arrays.push_back(arrow::MakeArray(arrow::ArrayData::Make(type, n, std::move(bufs), validity ? arrow::kUnknownNullCount : 0)));
    fields.push_back(arrow::field(col->name(), type));
  }
  if (cyl_stream_sync(stream) != CYL_OK) return Status(Code::ExecutionError, cyl_last_error());
  output = arrow::Table::Make(arrow::schema(fields), arrays);
  return Status::OK();
}

inline std::shared_ptr<arrow::Table> Table::get_table() {
  std::shared_ptr<arrow::Table> t;
  return ToArrowTable(t).is_ok() ? t : nullptr;
}

}  // namespace cylon
