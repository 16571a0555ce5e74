// cylon::Column -- one contiguous Arrow-layout array whose buffers live in HBM (validity bitmap, int32 offsets,
// data). It plays the role arrow::Array / cylon::Column play in the reference (cpp/src/cylon/column.hpp) for the
// hot path; buffers are owned (stream-ordered free) unless the column was made as a non-owning view.
#pragma once
#include <cstdint>
#include <memory>
#include <string>

#include "../cylon_b200_types.h"
#include "status.hpp"

namespace cylon {

class Column {
 public:
  Column() = default;
  ~Column();
  Column(const Column &) = delete;
  Column &operator=(const Column &) = delete;

  // allocate an uninitialised device column: `length` rows, `data_bytes` bytes of data, optional validity/offsets
  static Status Allocate(const std::string &name, int32_t type, int32_t byte_width, int64_t length, int64_t data_bytes,
                         bool with_validity, void *stream, std::shared_ptr<Column> *out);
  // copy host buffers (Arrow layout) to the device
  static Status FromHost(const std::string &name, const cyl_column_t &host, void *stream, std::shared_ptr<Column> *out);
  // wrap device buffers owned by someone else (they must outlive the column)
  static std::shared_ptr<Column> View(const std::string &name, const cyl_column_t &device_buffers);

  const std::string &name() const { return name_; }
  void set_name(const std::string &n) { name_ = n; }
  int32_t type() const { return desc_.type; }
  int32_t byte_width() const { return desc_.byte_width; }
  int64_t length() const { return desc_.length; }
  bool is_var_width() const { return desc_.type == CYL_STRING || desc_.type == CYL_BINARY; }
  bool has_validity() const { return desc_.validity != nullptr; }
  int64_t data_bytes() const { return data_bytes_; }
  const cyl_column_t &desc() const { return desc_; }  // device pointers
  uint8_t *mutable_validity() { return const_cast<uint8_t *>(desc_.validity); }
  int32_t *mutable_offsets() { return const_cast<int32_t *>(desc_.offsets); }
  void *mutable_data() { return const_cast<void *>(desc_.data); }
  void drop_validity();  // frees the bitmap (all rows valid)
  // fixed-width columns only: keep the first `rows` rows of a buffer that was allocated with spare capacity
  void truncate(int64_t rows) {
    desc_.length = rows;
    data_bytes_ = rows * (int64_t)desc_.byte_width;
  }

  // device -> host copies into caller buffers sized by length()/data_bytes(); any pointer may be null to skip
  Status ToHost(void *data, int32_t *offsets, uint8_t *validity, void *stream) const;

 private:
  std::string name_;
  cyl_column_t desc_{};
  int64_t data_bytes_ = 0;
  bool owns_ = false;
  void *stream_ = nullptr;
};

}  // namespace cylon
