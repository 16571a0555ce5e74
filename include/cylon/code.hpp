// cylon::Code -- numerically identical to the reference's enum (cpp/src/cylon/code.cpp:18-39), which mirrors
// arrow::StatusCode. The C-ABI returns the same values as plain ints (include/cylon_b200_types.h: cyl_code_t).
#pragma once
namespace cylon {
enum Code {
  OK = 0,
  OutOfMemory = 1,
  KeyError = 2,
  TypeError = 3,
  Invalid = 4,
  IOError = 5,
  CapacityError = 6,
  IndexError = 7,
  UnknownError = 9,
  NotImplemented = 10,
  SerializationError = 11,
  GpuMemoryError = 12,
  RError = 13,
  CodeGenError = 40,
  ExpressionValidationError = 41,
  ExecutionError = 42,
  AlreadyExists = 45
};
}  // namespace cylon
