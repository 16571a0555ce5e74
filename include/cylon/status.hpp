// cylon::Status -- same shape as the reference's (cpp/src/cylon/status.hpp:21-63): an int code plus a message,
// is_ok(), get_code(), get_msg(), Status::OK().
#pragma once
#include <string>
#include <utility>

#include "code.hpp"

namespace cylon {
class Status {
 public:
  Status() = default;
  Status(int code, std::string msg) : code_(code), msg_(std::move(msg)) {}
  explicit Status(int code) : code_(code) {}
  explicit Status(Code code) : code_(code) {}
  Status(Code code, std::string msg) : code_(code), msg_(std::move(msg)) {}
  int get_code() const { return code_; }
  bool is_ok() const { return code_ == Code::OK; }
  static Status OK() { return Status(Code::OK); }
  const std::string &get_msg() const { return msg_; }

 private:
  int code_ = Code::OK;
  std::string msg_;
};
}  // namespace cylon

#define RETURN_CYLON_STATUS_IF_FAILED(expr) \
  do {                                      \
    const auto &_st = (expr);               \
    if (!_st.is_ok()) return _st;           \
  } while (0)
