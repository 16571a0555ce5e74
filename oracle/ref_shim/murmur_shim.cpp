// extern "C" door onto the reference's own MurmurHash3_x86_32 (cpp/src/cylon/util/murmur3.cpp:76-117),
// compiled from the sources where they lie under /root/reference (never copied). Checker only.
#include <cstdint>
#include "util/murmur3.hpp"

extern "C" uint32_t ref_murmur3_x86_32(const void *key, int len, uint32_t seed) {
  uint32_t out = 0;
  cylon::util::MurmurHash3_x86_32(key, len, seed, &out);
  return out;
}
