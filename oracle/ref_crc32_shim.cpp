// TEST INFRASTRUCTURE ONLY.  extern "C" door to the reference's own calc_crc32, compiled from
// /root/reference/jubatus/server/common/crc32.cpp where it lies (oracle/Makefile target `ref`,
// output oracle/_ref/libref_crc32.so).  Used by tests/ to check the model-file container's
// checksum (jubatus_b200/host/model_file.hpp) against the reference implementation itself.
#include <stddef.h>
#include <stdint.h>

#include "crc32.hpp"   // -I/root/reference/jubatus/server/common

extern "C" uint32_t ref_calc_crc32(const char* data, size_t size, uint32_t crc) {
  return jubatus::server::common::calc_crc32(data, size, crc);
}
