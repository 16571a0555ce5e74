/*
 * cylon_b200_types.h -- plain-C types shared by the CUDA library's C-ABI
 * (cylon_b200.h), the C++ host layer (include/cylon/) and the CPU oracle
 * (oracle/, test infrastructure only).
 *
 * A column is described the way Arrow's C data interface lays a primitive or
 * binary array out: an optional validity bitmap (LSB-first, 1 = valid), an
 * optional int32 offsets buffer (variable-width types) and a data buffer.
 * For the CUDA library every pointer is a DEVICE pointer; for the oracle every
 * pointer is a host pointer.
 *
 * Type ids are numerically equal to arrow::Type::type, which is what the
 * reference switches on (cpp/src/cylon/arrow/arrow_partition_kernels.cpp:307-330,
 * cpp/src/cylon/arrow/arrow_kernels.cpp:172-195).
 */
#ifndef CYLON_B200_TYPES_H_
#define CYLON_B200_TYPES_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum cyl_type {
  CYL_BOOL = 1, /* bit-packed in Arrow; not supported on the path (no splitter in the reference either) */
  CYL_UINT8 = 2,
  CYL_INT8 = 3,
  CYL_UINT16 = 4,
  CYL_INT16 = 5,
  CYL_UINT32 = 6,
  CYL_INT32 = 7,
  CYL_UINT64 = 8,
  CYL_INT64 = 9,
  CYL_HALF_FLOAT = 10, /* not supported (no hash kernel in the reference) */
  CYL_FLOAT = 11,
  CYL_DOUBLE = 12,
  CYL_STRING = 13,
  CYL_BINARY = 14,
  CYL_FIXED_SIZE_BINARY = 15,
  CYL_DATE32 = 16,
  CYL_DATE64 = 17,
  CYL_TIMESTAMP = 18,
  CYL_TIME32 = 19,
  CYL_TIME64 = 20
} cyl_type_t;

/* One contiguous (single-chunk) Arrow array. */
typedef struct cyl_column {
  int32_t type;            /* cyl_type_t */
  int32_t byte_width;      /* bytes per value for fixed-width types; 0 for STRING/BINARY */
  int64_t length;          /* rows */
  int64_t null_count;      /* 0 when validity == NULL */
  const uint8_t *validity; /* bitmap, ceil(length/8) bytes, or NULL (= all valid) */
  const int32_t *offsets;  /* STRING/BINARY: length+1 int32 offsets into data; else NULL */
  const void *data;        /* values (fixed width) or bytes (variable width) */
} cyl_column_t;

/* cylon::Code (cpp/src/cylon/code.cpp:18-39), numerically equal to arrow::StatusCode. */
typedef enum cyl_code {
  CYL_OK = 0,
  CYL_OutOfMemory = 1,
  CYL_KeyError = 2,
  CYL_TypeError = 3,
  CYL_Invalid = 4,
  CYL_IOError = 5,
  CYL_CapacityError = 6,
  CYL_IndexError = 7,
  CYL_UnknownError = 9,
  CYL_NotImplemented = 10,
  CYL_SerializationError = 11,
  CYL_GpuMemoryError = 12,
  CYL_ExecutionError = 42,
  CYL_AlreadyExists = 45
} cyl_code_t;

/* cylon::join::config::JoinType / JoinAlgorithm (cpp/src/cylon/join/join_config.hpp:26-31). */
typedef enum cyl_join_type { CYL_JOIN_INNER = 0, CYL_JOIN_LEFT = 1, CYL_JOIN_RIGHT = 2, CYL_JOIN_FULL_OUTER = 3 } cyl_join_type_t;
typedef enum cyl_join_algorithm { CYL_JOIN_SORT = 0, CYL_JOIN_HASH = 1 } cyl_join_algorithm_t;

/* cylon::compute::AggregationOpId (cpp/src/cylon/compute/aggregate_kernels.hpp:40-50). */
typedef enum cyl_agg_op {
  CYL_AGG_SUM = 0, CYL_AGG_MIN = 1, CYL_AGG_MAX = 2, CYL_AGG_COUNT = 3,
  CYL_AGG_MEAN = 4, CYL_AGG_VAR = 5, CYL_AGG_STDDEV = 8 /* 6 = NUNIQUE, 7 = QUANTILE: not on the GPU path */
} cyl_agg_op_t;

/* Row-set operations of cpp/src/cylon/table.cpp:531-721,923-982. */
typedef enum cyl_setop { CYL_SETOP_UNION = 0, CYL_SETOP_SUBTRACT = 1, CYL_SETOP_INTERSECT = 2, CYL_SETOP_UNIQUE = 3 } cyl_setop_t;

#ifdef __cplusplus
}
#endif
#endif /* CYLON_B200_TYPES_H_ */
