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

#include <type_traits>

#if defined(__CUDACC__)
#define B200P_HD __host__ __device__ __forceinline__
#define B200P_D __device__ __forceinline__
#else
#define B200P_HD inline
#define B200P_D inline
#endif

namespace b200p {

#if defined(__CUDA_ARCH__)
B200P_D uint32_t Load4(const uint8_t *p) { return __ldg(reinterpret_cast<const uint32_t *>(p)); }
/* x << s for s = 0..32 (32 gives 0) */
B200P_D uint32_t ShlClamp(uint32_t x, uint32_t s) { return __funnelshift_lc(0u, x, s); }
/* bytes k.. of the 8-byte pair (hi:lo), k = 0..3 */
B200P_D uint32_t BytesFrom(uint32_t lo, uint32_t hi, uint32_t k) { return __funnelshift_r(lo, hi, k * 8u); }
B200P_D int FirstSet(uint32_t x) { return __ffs(static_cast<int>(x)) - 1; }
B200P_D float U32ToFloat(uint32_t x) { return __uint2float_rn(x); }
B200P_D float Fraction8(uint32_t f8) {
  return __double2float_rn(__dmul_rn(__uint2double_rn(f8), 1e-8));
}
B200P_D float Fraction9(uint32_t f9) {
  return __double2float_rn(__dmul_rn(__uint2double_rn(f9), 1e-9));
}
#else
inline uint32_t Load4(const uint8_t *p) {
  uint32_t a = 0;
  for (int i = 0; i < 4; ++i) a |= static_cast<uint32_t>(p[i]) << (8 * i);
  return a;
}
inline uint32_t ShlClamp(uint32_t x, uint32_t s) { return s >= 32u ? 0u : (x << s); }
inline uint32_t BytesFrom(uint32_t lo, uint32_t hi, uint32_t k) {
  return k == 0u ? lo : ((lo >> (8u * k)) | (hi << (32u - 8u * k)));
}
inline int FirstSet(uint32_t x) { return __builtin_ffs(static_cast<int>(x)) - 1; }
inline float U32ToFloat(uint32_t x) { return static_cast<float>(x); }
inline float Fraction8(uint32_t f8) {
  volatile double q = static_cast<double>(f8) * 1e-8;
  return static_cast<float>(q);
}
inline float Fraction9(uint32_t f9) {
  volatile double q = static_cast<double>(f9) * 1e-9;
  return static_cast<float>(q);
}
#endif


/* 16 bytes of text starting at some position, as four little-endian words */
struct Window {
  uint32_t x0, x1, x2, x3;
};

/* A chunk of text in HBM. Reads at or beyond `size` yield NUL, which is what the reference
   sees past a chunk (it zero-pads, reference src/io/input_split_base.cc:247).
   The converters below are templates over the text source; a source offers
     at(i)          one byte (NUL past the end)
     Window16(p,w)  the 16 bytes at p as four words, false when they cannot be had cheaply
     Tail8(p,a,b)   bytes 16..23 after p as two words, same convention
   Text is the plain one (global memory); the fused kernel adds a source that serves windows
   from the tile staged in shared memory (tile_kernels.cuh). */
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
  /* false if the aligned loads would touch memory past the chunk
     ([p & ~3, (p & ~3) + 20) must be readable) */
  B200P_HD bool Window16(uint32_t p, Window *w) const {
    const uint32_t a = p & ~3u;
    if (a + 20u > size) return false;
    const uint32_t w0 = Load4(data + a), w1 = Load4(data + a + 4), w2 = Load4(data + a + 8),
                   w3 = Load4(data + a + 12), w4 = Load4(data + a + 16);
    const uint32_t k = p & 3u;
    w->x0 = BytesFrom(w0, w1, k);
    w->x1 = BytesFrom(w1, w2, k);
    w->x2 = BytesFrom(w2, w3, k);
    w->x3 = BytesFrom(w3, w4, k);
    return true;
  }
  B200P_HD bool Tail8(uint32_t p, uint32_t *x4, uint32_t *x5) const {
    const uint32_t a = p & ~3u, k = p & 3u;
    if (a + 28u > size) return false;
    const uint32_t w4 = Load4(data + a + 16), w5 = Load4(data + a + 20), w6 = Load4(data + a + 24);
    *x4 = BytesFrom(w4, w5, k);
    *x5 = BytesFrom(w5, w6, k);
    return true;
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
template <typename TXT>
B200P_HD float ToFloat(const TXT &t, uint32_t p, uint32_t *end) {
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
template <bool WIDE, typename TXT>
B200P_HD uint64_t ToUint(const TXT &t, uint32_t p, bool *negative) {
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
template <typename TXT>
B200P_HD uint64_t ToQid(const TXT &t, uint32_t p) {
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


/* ------------------------------------------------------------------------------------------
 * Register-window fast paths.
 *
 * The general converters above fetch one byte per loop trip. Almost every real token is short
 * and plain ("-7.60915", "0.137604", "52172", "1"), so the fast paths pull 16 bytes of the
 * token into registers with three aligned 8-byte loads, find the digit runs with SWAR
 * arithmetic, and convert from registers.  Anything they cannot prove identical to the general
 * code -- leading whitespace, > 7 integer digits (fp32 accumulation starts rounding at 2^24),
 * > 7 fraction digits, an exponent, a token that does not end inside the window, a window
 * that would read past the chunk -- is handed to the general code.
 *
 * Fraction: the reference computes float(double(digits) / double(10^k)).  For k <= 9 and every
 * digits < 10^k, float(double(digits) * RN(10^-k)) is the same fp32 value; that was checked
 * exhaustively on the CPU (all 1 111 111 111 cases), so one DMUL replaces the
 * IEEE fp64 division.
 * ------------------------------------------------------------------------------------------ */
/* bit 7 of every byte that is NOT an ASCII digit */
B200P_HD uint32_t NonDigitFlags(uint32_t x) {
  const uint32_t y = x ^ 0x30303030u;
  return (((y & 0x7f7f7f7fu) + 0x76767676u) | y) & 0x80808080u;
}

/* byte i (0..7) of the pair (hi:lo) */
B200P_HD uint32_t ByteOf(uint32_t lo, uint32_t hi, uint32_t i) {
  return ((i < 4u ? lo : hi) >> ((i & 3u) * 8u)) & 0xffu;
}

/* four digit values d0 d1 d2 d3 (one per byte, d0 in the low byte = most significant) -> number,
   added to `plus`.  z * 10 + (z >> 8) leaves d0d1 in byte 0 and d2d3 in byte 2 (bytes 1 and 3 are
   by-products); a byte dot product with weights (100, 0, 1, 0) finishes the job in one
   instruction. */
B200P_HD uint32_t Pack4(uint32_t z, uint32_t plus) {
  const uint32_t t = z * 10u + (z >> 8);
#if defined(__CUDA_ARCH__)
  return __dp4a(t, 0x00010064u, plus);
#else
  return (t & 0xffu) * 100u + ((t >> 16) & 0xffu) + plus;
#endif
}

template <typename TXT>
B200P_HD bool ToFloatWindow(const TXT &t, uint32_t p, float *out) {
  Window w;
  if (!t.Window16(p, &w)) return false;
  const uint32_t c0 = w.x0 & 0xffu;
  if (c0 <= 0x20u) return false;           /* whitespace / control first: general code */
  const bool negative = c0 == '-';
  const uint32_t s0 = (negative || c0 == '+') ? 1u : 0u;   /* sign bytes */
  /* integer digits: at most 7, so that the reference's fp32 accumulation is exact (< 2^24) */
  const uint32_t nd0 = NonDigitFlags(w.x0) & (s0 != 0u ? 0xffffff00u : 0xffffffffu);
  uint32_t term, ip, f0; /* f0 = window offset of the byte after the terminator of the digits */
  if (nd0 != 0u) {
    /* sign + digits end inside the first word: drop the sign byte, left-pad the digits to four
       places and pack */
    const uint32_t bit = static_cast<uint32_t>(FirstSet(nd0)) & ~7u;   /* 8 * (s0 + n1) */
    f0 = (bit >> 3) + 1u;
    term = (w.x0 >> bit) & 0xffu;
    const uint32_t digits = w.x0 & (s0 != 0u ? 0x0f0f0f00u : 0x0f0f0f0fu);
    ip = Pack4(ShlClamp(digits, 32u - bit), 0u);
  } else {
    const uint32_t nd1 = NonDigitFlags(w.x1);
    if (nd1 == 0u) return false;            /* digits run past the second word */
    const uint32_t bit = static_cast<uint32_t>(FirstSet(nd1)) & ~7u;
    const uint32_t n = 4u + (bit >> 3);     /* s0 + n1 = 4..7 */
    f0 = n + 1u;
    term = (w.x1 >> bit) & 0xffu;
    /* the first word packed (sign byte as a leading zero), the rest appended one by one */
    ip = Pack4(w.x0 & (s0 != 0u ? 0x0f0f0f00u : 0x0f0f0f0fu), 0u);
    const uint32_t y1 = w.x1 & 0x0f0f0f0fu;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (uint32_t i = 0; i < 3u; ++i) {
      if (4u + i < n) ip = ip * 10u + ((y1 >> (8u * i)) & 0xffu);
    }
  }
  float v = U32ToFloat(ip);
  if (term == '.') {
    /* fraction: the 8 bytes after the dot, digits beyond the run zeroed = the digits right-padded
       to 8 places; float(double(F8) * 1e-8) == float(double(f) / double(10^k)) for every k <= 8
       (exhaustive CPU check) */
    uint32_t a, b, c;                        /* f0 = 1 .. 8 */
    if (f0 < 4u) {
      a = w.x0; b = w.x1; c = w.x2;
    } else if (f0 < 8u) {
      a = w.x1; b = w.x2; c = w.x3;
    } else {
      a = w.x2; b = w.x3; c = 0u;
    }
    const uint32_t r0 = BytesFrom(a, b, f0 & 3u);
    const uint32_t ndr0 = NonDigitFlags(r0);
    uint32_t n2, f8;
    if (ndr0 != 0u) {
      const uint32_t bit = static_cast<uint32_t>(FirstSet(ndr0)) & ~7u;
      n2 = bit >> 3;
      term = (r0 >> bit) & 0xffu;
      f8 = Pack4(r0 & 0x0f0f0f0fu & ((1u << bit) - 1u), 0u) * 10000u;
    } else {
      const uint32_t r1 = BytesFrom(b, c, f0 & 3u);
      const uint32_t ndr1 = NonDigitFlags(r1);
      const uint32_t high = Pack4(r0 & 0x0f0f0f0fu, 0u) * 10000u;
      if (ndr1 != 0u) {
        const uint32_t bit = static_cast<uint32_t>(FirstSet(ndr1)) & ~7u;
        n2 = 4u + (bit >> 3);
        term = (r1 >> bit) & 0xffu;
        f8 = Pack4(r1 & 0x0f0f0f0fu & ((1u << bit) - 1u), high);
      } else {
        /* exactly eight digits ("%.6g" of 0.00123456): the terminator is the next byte, which is
           byte (f0 & 3) of the third word (the f0 + n2 test below rejects the case where that
           word lies outside the window) */
        n2 = 8u;
        const uint32_t k = f0 & 3u;
        term = (c >> (8u * k)) & 0xffu;
        f8 = Pack4(r1 & 0x0f0f0f0fu, high);
        if (term - '0' < 10u) {
          /* a ninth digit ("%.6g" of 0.000123456): F9 * 1e-9 rounds like F9 / 10^9 for every
             F9 < 10^9 (exhaustive CPU check); rare, so it finishes on its own */
          if (k == 3u || f0 > 6u) return false;            /* terminator outside the three words */
          const uint32_t after = (c >> (8u * k + 8u)) & 0xffu;
          if (after - '0' < 10u || (after | 0x20u) == 'e') return false;   /* ten digits / exponent */
          v = AddRn(v, Fraction9(f8 * 10u + (term - '0')));
          *out = negative ? -v : v;
          return true;
        }
      }
    }
    if (f0 + n2 > 15u) return false;         /* the terminator must be a real byte of the window */
    v = AddRn(v, Fraction8(f8));
  }
  if ((term | 0x20u) == 'e') return false;   /* exponent: general code */
  *out = negative ? -v : v;
  return true;
}

/* float token at p: window fast path, else the general converter */
template <typename TXT>
B200P_HD float ToFloatFast(const TXT &t, uint32_t p) {
  float v;
  if (ToFloatWindow(t, p, &v)) return v;
  return ToFloat(t, p, 0);
}

/* 10^m for m = 0..3, out of one 64-bit constant */
B200P_HD uint32_t TenTo(uint32_t m) {
  return static_cast<uint32_t>(0x03e80064000a0001ull >> (16u * m)) & 0xffffu;
}

/* the first m (0..3) digit bytes of x as a number */
B200P_HD uint32_t PackLeading(uint32_t x, uint32_t m) {
  return Pack4(ShlClamp(x & 0x0f0f0f0fu, 32u - 8u * m), 0u);
}

/*
 * Unsigned token at p: 1..23 digits followed by a non-digit, four digits per step (Pack4) and
 * one multiplication per word instead of one per digit.  All arithmetic is in the ring the
 * reference accumulates in (mod 2^32 through Str2T<uint32_t>, mod 2^64 for uint64_t,
 * strtonum.h:103-190), so overlong numbers wrap to the same value.  Tokens of 16 digits and more
 * (64-bit hashes) read two more words; 24 and more, a sign or a blank in front go to ToUint.
 */
template <bool WIDE, typename TXT>
B200P_HD bool ToUintWindow(const TXT &t, uint32_t p, uint64_t *out) {
  Window w;
  if (!t.Window16(p, &w)) return false;
  typedef typename std::conditional<WIDE, uint64_t, uint32_t>::type Acc;
  uint32_t flags = NonDigitFlags(w.x0);
  if (flags != 0u) {
    const uint32_t m = static_cast<uint32_t>(FirstSet(flags)) >> 3;
    if (m == 0u) return false;              /* sign, whitespace, empty: general code */
    *out = PackLeading(w.x0, m);
    return true;
  }
  uint32_t low = Pack4(w.x0 & 0x0f0f0f0fu, 0u);                  /* digits 0..3 */
  flags = NonDigitFlags(w.x1);
  if (flags != 0u) {
    const uint32_t m = static_cast<uint32_t>(FirstSet(flags)) >> 3;
    *out = low * TenTo(m) + PackLeading(w.x1, m);                 /* < 10^7 */
    return true;
  }
  low = low * 10000u + Pack4(w.x1 & 0x0f0f0f0fu, 0u);             /* digits 0..7, < 10^8 */
  flags = NonDigitFlags(w.x2);
  if (flags != 0u) {
    const uint32_t m = static_cast<uint32_t>(FirstSet(flags)) >> 3;
    *out = static_cast<Acc>(static_cast<Acc>(low) * TenTo(m) + PackLeading(w.x2, m));
    return true;
  }
  Acc v = static_cast<Acc>(static_cast<Acc>(low) * 10000u + Pack4(w.x2 & 0x0f0f0f0fu, 0u));   /* 12 digits */
  flags = NonDigitFlags(w.x3);
  if (flags != 0u) {
    const uint32_t m = static_cast<uint32_t>(FirstSet(flags)) >> 3;
    *out = static_cast<Acc>(v * TenTo(m) + PackLeading(w.x3, m));
    return true;
  }
  /* 16 digits and more: bytes 16..23 of the token */
  uint32_t x4, x5;
  if (!t.Tail8(p, &x4, &x5)) return false;
  v = static_cast<Acc>(v * 10000u + Pack4(w.x3 & 0x0f0f0f0fu, 0u));                            /* 16 digits */
  flags = NonDigitFlags(x4);
  if (flags != 0u) {
    const uint32_t m = static_cast<uint32_t>(FirstSet(flags)) >> 3;
    *out = static_cast<Acc>(v * TenTo(m) + PackLeading(x4, m));
    return true;
  }
  v = static_cast<Acc>(v * 10000u + Pack4(x4 & 0x0f0f0f0fu, 0u));                              /* 20 digits */
  flags = NonDigitFlags(x5);
  if (flags == 0u) return false;            /* 24 digits and more */
  const uint32_t m = static_cast<uint32_t>(FirstSet(flags)) >> 3;
  *out = static_cast<Acc>(v * TenTo(m) + PackLeading(x5, m));
  return true;
}

template <bool WIDE, typename TXT>
B200P_HD uint64_t ToUintFast(const TXT &t, uint32_t p, bool *negative) {
  uint64_t v;
  if (ToUintWindow<WIDE>(t, p, &v)) return v;
  return ToUint<WIDE>(t, p, negative);
}

}  // namespace b200p
#endif  // DMLC_CORE_B200_STRTONUM_DEVICE_H_
