/*
 * Per-token numeric conversion for the sm_100a text parsers.
 *
 * Bit-exact device counterparts of the reference converters:
 *   to_float  <-> dmlc::data::strtof          reference src/data/strtonum.h:37-97
 *   to_uint   <-> Str2T<uint32_t/uint64_t>    reference src/data/strtonum.h:126-150,186-207
 *   to_qid    <-> ::atoll (glibc)             reference src/data/libsvm_parser.h:69-75
 *
 * Every floating-point step uses an explicitly rounded intrinsic (__fmul_rn / __fadd_rn /
 * __fdiv_rn / __ddiv_rn ...) so nvcc can never contract `acc * 10 + d` into an FMA: the
 * reference build has no FMA (-msse2 only, reference Makefile:16,24-26) and rounds twice.
 *
 * The header also compiles as plain C++ (B200P_HOST_EMULATION) so the not-gpu tests can run the
 * very same token code on the CPU against the oracle; that build is test-only and is not
 * linked into the product library.
 */
#ifndef DMLC_CORE_B200_STRTONUM_DEVICE_H_
#define DMLC_CORE_B200_STRTONUM_DEVICE_H_

#include <stdint.h>

#if defined(__CUDACC__)
#define B200P_HD __host__ __device__ __forceinline__
#define B200P_D __device__ __forceinline__
#else
#define B200P_HD inline
#define B200P_D inline
#endif

namespace b200p {

/* A chunk of text in HBM. Reads at or beyond `size` yield NUL, which is what the reference
   sees past a chunk (it zero-pads, reference src/io/input_split_base.cc:247). */
struct Text {
  const uint8_t *data;
  uint32_t size;
  B200P_HD uint32_t at(uint32_t i) const {
#if defined(__CUDA_ARCH__)
    return i < size ? static_cast<uint32_t>(__ldg(data + i)) : 0u;
#else
    return i < size ? static_cast<uint32_t>(data[i]) : 0u;
#endif
  }
};

/* character classes, reference src/data/strtonum.h:14-31 */
B200P_HD bool IsSpace(uint32_t c) {
  /* ' ' \t \n \f \r  (note: not \v) */
  return c == ' ' || c == '\t' || c == '\n' || c == '\f' || c == '\r';
}
B200P_HD bool IsBlank(uint32_t c) { return c == ' ' || c == '\t'; }
B200P_HD bool IsEol(uint32_t c) { return c == '\n' || c == '\r'; }
/* digitchar = 0-9 + - . e E, as one 64-bit membership test on (c - '+') */
B200P_HD bool IsDigitChar(uint32_t c) {
  const uint32_t t = c - 0x2Bu;
  return t < 59u && ((0x0400000004007FEDull >> t) & 1ull) != 0ull;
}

#if defined(__CUDA_ARCH__)
B200P_D float MulRn(float a, float b) { return __fmul_rn(a, b); }
B200P_D float AddRn(float a, float b) { return __fadd_rn(a, b); }
B200P_D float DivRn(float a, float b) { return __fdiv_rn(a, b); }
B200P_D float FractionRn(uint64_t digits, uint64_t pow10) {
  return __double2float_rn(__ddiv_rn(__ull2double_rn(digits), __ull2double_rn(pow10)));
}
#else
/* host emulation for the CPU unit test; built with -ffp-contract=off */
inline float MulRn(float a, float b) { volatile float r = a * b; return r; }
inline float AddRn(float a, float b) { volatile float r = a + b; return r; }
inline float DivRn(float a, float b) { volatile float r = a / b; return r; }
inline float FractionRn(uint64_t digits, uint64_t pow10) {
  volatile double q = static_cast<double>(digits) / static_cast<double>(pow10);
  return static_cast<float>(q);
}
#endif

B200P_HD float BitsToFloat(uint32_t b) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(b);
#else
  union { uint32_t u; float f; } x; x.u = b; return x.f;
#endif
}

/*
 * fp32 value of 10^e exactly as the reference builds it: `float scale = 1` multiplied by 1e8
 * (in fp64, rounded back to fp32) while e >= 8, then by 10.0 the same way
 * (reference src/data/strtonum.h:74,89-90).  Regenerated from oracle/_ref; see SURVEY.md App. A.
 */
B200P_HD float ScaleOfTen(uint32_t e) {
  switch (e) {
    case 0: return BitsToFloat(0x3f800000u); case 1: return BitsToFloat(0x41200000u);
    case 2: return BitsToFloat(0x42c80000u); case 3: return BitsToFloat(0x447a0000u);
    case 4: return BitsToFloat(0x461c4000u); case 5: return BitsToFloat(0x47c35000u);
    case 6: return BitsToFloat(0x49742400u); case 7: return BitsToFloat(0x4b189680u);
    case 8: return BitsToFloat(0x4cbebc20u); case 9: return BitsToFloat(0x4e6e6b28u);
    case 10: return BitsToFloat(0x501502f9u); case 11: return BitsToFloat(0x51ba43b7u);
    case 12: return BitsToFloat(0x5368d4a5u); case 13: return BitsToFloat(0x551184e7u);
    case 14: return BitsToFloat(0x56b5e621u); case 15: return BitsToFloat(0x58635fa9u);
    case 16: return BitsToFloat(0x5a0e1bcau); case 17: return BitsToFloat(0x5bb1a2bcu);
    case 18: return BitsToFloat(0x5d5e0b6bu); case 19: return BitsToFloat(0x5f0ac723u);
    case 20: return BitsToFloat(0x60ad78ecu); case 21: return BitsToFloat(0x6258d727u);
    case 22: return BitsToFloat(0x64078678u); case 23: return BitsToFloat(0x65a96816u);
    case 24: return BitsToFloat(0x6753c21cu); case 25: return BitsToFloat(0x69045952u);
    case 26: return BitsToFloat(0x6aa56fa6u); case 27: return BitsToFloat(0x6c4ecb90u);
    case 28: return BitsToFloat(0x6e013f3au); case 29: return BitsToFloat(0x6fa18f08u);
    case 30: return BitsToFloat(0x7149f2cau); case 31: return BitsToFloat(0x72fc6f7cu);
    case 32: return BitsToFloat(0x749dc5aeu); case 33: return BitsToFloat(0x7645371au);
    case 34: return BitsToFloat(0x77f684e0u); case 35: return BitsToFloat(0x799a130cu);
    case 36: return BitsToFloat(0x7b4097cfu); case 37: return BitsToFloat(0x7cf0bdc3u);
    default: return BitsToFloat(0x7e96769au);
  }
}

/*
 * Float token at position p.  Grammar (reference strtonum.h:37-97):
 *   [space]* [+-]? digit* ( '.' digit* )? ( [eE] [+-]? digit* )?
 * `end` (optional) receives the position of the first byte not consumed.
 */
B200P_HD float ToFloat(const Text &t, uint32_t p, uint32_t *end) {
  uint32_t c = t.at(p);
  while (IsSpace(c)) c = t.at(++p);
  bool negative = false;
  if (c == '-') {
    negative = true;
    c = t.at(++p);
  } else if (c == '+') {
    c = t.at(++p);
  }
  /* integer digits: fp32 accumulate, two roundings per digit (:52-54) */
  float acc = 0.0f;
  for (uint32_t d = c - '0'; d < 10u; d = c - '0') {
    acc = AddRn(MulRn(acc, 10.0f), static_cast<float>(static_cast<int>(d)));
    c = t.at(++p);
  }
  bool nan = false;
  if (c == '.') {
    /* fraction digits: two wrapping u64, fp64 divide, fp32 add (:57-68) */
    uint64_t digits = 0, pow10 = 1;
    c = t.at(++p);
    for (uint32_t d = c - '0'; d < 10u; d = c - '0') {
      digits = digits * 10ull + d;
      pow10 *= 10ull;
      c = t.at(++p);
    }
    if (pow10 == 0ull && digits == 0ull) {
      /* >= 64 fraction digits wrap pow10 to 0: 0/0.  SSE2 yields the default NaN
         (sign set, 0xffc00000 as fp32) and every later op propagates it unchanged;
         CUDA would canonicalise to 0x7fffffff, so it is carried as a flag instead. */
      nan = true;
    } else {
      acc = AddRn(acc, FractionRn(digits, pow10));
    }
  }
  if (c == 'e' || c == 'E') {
    c = t.at(++p);
    bool divide = false;
    if (c == '-') {
      divide = true;
      c = t.at(++p);
    } else if (c == '+') {
      c = t.at(++p);
    }
    uint32_t ex = 0; /* wraps mod 2^32 like the reference's `unsigned` (:84-86) */
    for (uint32_t d = c - '0'; d < 10u; d = c - '0') {
      ex = ex * 10u + d;
      c = t.at(++p);
    }
    if (ex > 38u) ex = 38u;
    const float scale = ScaleOfTen(ex);
    acc = divide ? DivRn(acc, scale) : MulRn(acc, scale);
  }
  if (end != 0) *end = p;
  if (nan) return BitsToFloat(negative ? 0x7fc00000u : 0xffc00000u);
  return negative ? -acc : acc;
}

/*
 * Unsigned index / field token at p.  WIDE=false mirrors Str2T<uint32_t> (accumulates in a
 * 32-bit int => result mod 2^32, strtonum.h:186-190); WIDE=true mirrors Str2T<uint64_t>.
 * A '-' sign makes the reference CHECK fail (:141): reported through *negative.
 */
template <bool WIDE>
B200P_HD uint64_t ToUint(const Text &t, uint32_t p, bool *negative) {
  uint32_t c = t.at(p);
  while (IsSpace(c)) c = t.at(++p);
  if (c == '-') {
    *negative = true;
    return 0;
  }
  if (c == '+') c = t.at(++p);
  if (WIDE) {
    uint64_t v = 0;
    for (uint32_t d = c - '0'; d < 10u; d = c - '0') {
      v = v * 10ull + d;
      c = t.at(++p);
    }
    return v;
  } else {
    uint32_t v = 0;
    for (uint32_t d = c - '0'; d < 10u; d = c - '0') {
      v = v * 10u + d;
      c = t.at(++p);
    }
    return v;
  }
}

/* qid: glibc atoll -- C-locale whitespace (' ' and \t..\r), sign, digits, saturating. */
B200P_HD uint64_t ToQid(const Text &t, uint32_t p) {
  uint32_t c = t.at(p);
  while (c == ' ' || (c - 9u) < 5u) c = t.at(++p);
  bool negative = false;
  if (c == '-') {
    negative = true;
    c = t.at(++p);
  } else if (c == '+') {
    c = t.at(++p);
  }
  const uint64_t cap = negative ? (1ull << 63) : ((1ull << 63) - 1ull);
  uint64_t v = 0;
  bool saturated = false;
  for (uint32_t d = c - '0'; d < 10u; d = c - '0') {
    if (!saturated) {
      if (v > (cap - d) / 10ull) {
        saturated = true;
        v = cap;
      } else {
        v = v * 10ull + d;
      }
    }
    c = t.at(++p);
  }
  return negative ? (0ull - v) : v;
}

}  // namespace b200p
#endif  // DMLC_CORE_B200_STRTONUM_DEVICE_H_
