// cylon::compute::AggregationOpId -- same enumerators and values as the reference
// (cpp/src/cylon/compute/aggregate_kernels.hpp:40-50). SUM / MIN / MAX / COUNT / MEAN / VAR / STDDEV run on the GPU; NUNIQUE and
// QUANTILE are outside the accelerated path and return Code::NotImplemented.
#pragma once
namespace cylon {
namespace compute {
enum AggregationOpId { SUM, MIN, MAX, COUNT, MEAN, VAR, NUNIQUE, QUANTILE, STDDEV };
// output column prefix, KernelTraits<op,T>::name() (aggregate_kernels.hpp:186-260)
inline const char *AggregationPrefix(AggregationOpId op) {
  switch (op) {
    case SUM: return "sum_";
    case MIN: return "min_";
    case MAX: return "max_";
    case COUNT: return "count_";
    case MEAN: return "mean_";
    case VAR: return "var_";
    case NUNIQUE: return "nunique_";
    case QUANTILE: return "quantile_";
    default: return "std_";
  }
}
}  // namespace compute
}  // namespace cylon
