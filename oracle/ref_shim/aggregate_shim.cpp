// TEST INFRASTRUCTURE ONLY. Compiles the reference's aggregation kernels (SumKernel / MinKernel / MaxKernel / CountKernel,
// /root/reference/cpp/src/cylon/compute/aggregate_kernels.hpp:410-485, included IN PLACE, nothing copied) and drives
// them the way the reference's group-by does (groupby/hash_groupby.cpp:142-201): default options, ONE initial state
// copied to every group, Update(value, state[group id]) per row in row order, Finalize per group. Exposed through a
// C function so that tests can pin the oracle's aggregate arithmetic (including MAX's numeric_limits<T>::min() start
// value and integer wrap-around) to the reference's own code.
#include <algorithm>
#include <cstdint>
#include <limits>
#include <memory>
#include <tuple>
#include <vector>

#include <compute/aggregate_kernels.hpp>

namespace {

using cylon::compute::AggregationOpId;

template <AggregationOpId OP, typename T>
int run(const void *values, const int64_t *gid, int64_t n, int64_t groups, void *out) {
  using Traits = cylon::compute::KernelTraits<OP, T>;
  using State = typename Traits::State;
  using Result = typename Traits::ResultT;
  auto kernel = cylon::compute::CreateAggregateKernel<OP, T>();
  typename Traits::Options opt;
  kernel->Setup(&opt);
  State first;
  kernel->InitializeState(&first);
  std::vector<State> states((size_t)groups, first);
  const T *v = static_cast<const T *>(values);
  for (int64_t i = 0; i < n; i++) {
    if (gid[i] < 0 || gid[i] >= groups) return 2;
    T x = v[i];
    kernel->Update(&x, &states[(size_t)gid[i]]);
  }
  Result *o = static_cast<Result *>(out);
  for (int64_t g = 0; g < groups; g++) kernel->Finalize(&states[(size_t)g], &o[g]);
  return 0;
}

template <typename T>
int by_op(int op, const void *values, const int64_t *gid, int64_t n, int64_t groups, void *out) {
  switch (op) {
    case cylon::compute::SUM: return run<cylon::compute::SUM, T>(values, gid, n, groups, out);
    case cylon::compute::MIN: return run<cylon::compute::MIN, T>(values, gid, n, groups, out);
    case cylon::compute::MAX: return run<cylon::compute::MAX, T>(values, gid, n, groups, out);
    case cylon::compute::COUNT: return run<cylon::compute::COUNT, T>(values, gid, n, groups, out);
    default: return 1;
  }
}

}  // namespace

// dtype: 0 i8, 1 i16, 2 i32, 3 i64, 4 u8, 5 u16, 6 u32, 7 u64, 8 f32, 9 f64. op: the reference's AggregationOpId value
// (SUM 0, MIN 1, MAX 2, COUNT 3). out: `groups` values of the kernel's ResultT (T, or int64 for COUNT).
extern "C" int ref_aggregate(int op, int dtype, const void *values, const int64_t *gid, int64_t n, int64_t groups, void *out) {
  switch (dtype) {
    case 0: return by_op<int8_t>(op, values, gid, n, groups, out);
    case 1: return by_op<int16_t>(op, values, gid, n, groups, out);
    case 2: return by_op<int32_t>(op, values, gid, n, groups, out);
    case 3: return by_op<int64_t>(op, values, gid, n, groups, out);
    case 4: return by_op<uint8_t>(op, values, gid, n, groups, out);
    case 5: return by_op<uint16_t>(op, values, gid, n, groups, out);
    case 6: return by_op<uint32_t>(op, values, gid, n, groups, out);
    case 7: return by_op<uint64_t>(op, values, gid, n, groups, out);
    case 8: return by_op<float>(op, values, gid, n, groups, out);
    case 9: return by_op<double>(op, values, gid, n, groups, out);
    default: return 1;
  }
}
