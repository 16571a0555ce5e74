// Force-included (-include) when compiling the reference IN PLACE against Arrow 24 (the reference pins Arrow 2.0).
// It restores two names Arrow 2 had and stands in for one reference header whose MemoryPool overrides no longer match
// Arrow's virtual signatures (ctx/arrow_memory_pool_utils.hpp, skipped through its include guard). Checker build only;
// the reference sources are not modified or copied.
#pragma once
#include <arrow/memory_pool.h>

#include <memory>
#include <string>
#include <string_view>

namespace arrow {
namespace util {
class string_view : public std::string_view {
 public:
  using std::string_view::string_view;
  string_view(std::string_view v) : std::string_view(v) {}  // NOLINT implicit on purpose
  std::string to_string() const { return std::string(*this); }
};
}  // namespace util
}  // namespace arrow
namespace std {
template <>
struct hash<arrow::util::string_view> {
  size_t operator()(const arrow::util::string_view &v) const noexcept { return std::hash<std::string_view>()(v); }
};
}  // namespace std

#include <arrow/type_fwd.h>
namespace arrow {
namespace internal {
// Arrow 2 had this in arrow/type.h; every type the shuffle carries except NA has a validity bitmap.
inline bool HasValidityBitmap(Type::type id) { return id != Type::NA; }
}  // namespace internal
}  // namespace arrow

#include <arrow/array/builder_primitive.h>
namespace arrow {
// util/copy_arrray.cpp:148 instantiates its numeric copy for BooleanType through NumericBuilder<BooleanType>, which
// Arrow 2 accepted. This specialisation only lets that one instantiation compile; the checker never builds BOOL columns.
template <>
class NumericBuilder<BooleanType> : public BooleanBuilder {
 public:
  using BooleanBuilder::BooleanBuilder;
};
}  // namespace arrow

#define CYLON_SRC_CYLON_CTX_ARROW_MEMORY_POOL_UTILS_HPP_
namespace cylon {
class CylonContext;
arrow::MemoryPool *ToArrowPool(const std::shared_ptr<cylon::CylonContext> &ctx);
}  // namespace cylon
