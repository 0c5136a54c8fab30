// TEST ONLY: host build of dmlc_core_b200/csrc/strtonum_device.h so the not-gpu tests can run
// the exact token-conversion code of the kernels on the CPU against the oracle.
#include <cstring>
#include "strtonum_device.h"

extern "C" {
uint32_t emul_strtof_bits(const char *s, int64_t *consumed) {
  b200p::Text t = {reinterpret_cast<const uint8_t *>(s), static_cast<uint32_t>(std::strlen(s))};
  uint32_t end = 0;
  float v = b200p::ToFloat(t, 0, &end);
  if (consumed) *consumed = end;
  uint32_t bits;
  std::memcpy(&bits, &v, 4);
  return bits;
}
uint64_t emul_strtouint(const char *s, int width, int *threw) {
  b200p::Text t = {reinterpret_cast<const uint8_t *>(s), static_cast<uint32_t>(std::strlen(s))};
  bool neg = false;
  uint64_t v = width == 8 ? b200p::ToUint<true>(t, 0, &neg) : b200p::ToUint<false>(t, 0, &neg);
  *threw = neg ? 1 : 0;
  return v;
}
int64_t emul_atoll(const char *s) {
  b200p::Text t = {reinterpret_cast<const uint8_t *>(s), static_cast<uint32_t>(std::strlen(s))};
  return static_cast<int64_t>(b200p::ToQid(t, 0));
}
// window fast paths: the token sits at `offset` inside a larger buffer (so every alignment of the
// 8-byte loads is exercised) followed by `tail` bytes of other text
uint32_t emul_fast_strtof_bits(const char *buf, uint32_t size, uint32_t offset, int *took_fast) {
  b200p::Text t = {reinterpret_cast<const uint8_t *>(buf), size};
  float w = 0.f;
  *took_fast = b200p::ToFloatWindow(t, offset, &w) ? 1 : 0;
  float v = b200p::ToFloatFast(t, offset);
  uint32_t bits;
  std::memcpy(&bits, &v, 4);
  return bits;
}
uint64_t emul_fast_strtouint(const char *buf, uint32_t size, uint32_t offset, int width, int *threw,
                             int *took_fast) {
  b200p::Text t = {reinterpret_cast<const uint8_t *>(buf), size};
  bool neg = false;
  uint64_t w = 0;
  *took_fast = (width == 8 ? b200p::ToUintWindow<true>(t, offset, &w)
                           : b200p::ToUintWindow<false>(t, offset, &w)) ? 1 : 0;
  uint64_t v = width == 8 ? b200p::ToUintFast<true>(t, offset, &neg)
                          : b200p::ToUintFast<false>(t, offset, &neg);
  *threw = neg ? 1 : 0;
  return v;
}
int emul_is_digitchar(unsigned c) { return b200p::IsDigitChar(c) ? 1 : 0; }
}
