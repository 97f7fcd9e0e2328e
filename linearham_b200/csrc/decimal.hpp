// Decimal text -> IEEE double, correctly rounded, for the device Newick parser (newick_dev.cu) and,
// compiled for the host, for its CPU test against strtod (tests/test_decimal.py).
//
// The reference reads branch lengths with atof (libpll parse_utree.y / lex_utree.l); the host parser of
// this repo with std::from_chars (host/newick.cpp). Both return the correctly rounded double, so any
// correctly rounding conversion gives the same bits. This is the Eisel-Lemire algorithm (D. Lemire,
// "Number parsing at a gigabyte per second", SPE 2021; N. Mushtak & D. Lemire, "Fast number parsing
// without fallback", SPE 2023): a decimal significand w < 2^64 and exponent q are turned into a double
// with one or two 64x64->128-bit multiplications by a truncated 128-bit power of five.
//
// parse_decimal() accepts exactly the plain form  [-]digits[.digits][(e|E)[+|-]digits]  (or .digits) and
// reports everything it does not want to decide itself - more than 19 significant digits, results outside
// the normal range, other spellings (inf, nan, hex, leading '+') - as kDecimalHard; the caller then takes
// the tree to the host parser (host/newick.cpp), which handles them as strtod does.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define LH_HD __host__ __device__ __forceinline__
#else
#define LH_HD inline
#endif

namespace lhb {

static const uint64_t kPow5_128_host[2 * 651] = {
#include "pow5_128.inc"
};
#if defined(__CUDACC__)
static __device__ const uint64_t kPow5_128_dev[2 * 651] = {
#include "pow5_128.inc"
};
#endif
#if defined(__CUDA_ARCH__)
#define LH_POW5(i) __ldg(&kPow5_128_dev[i])
#else
#define LH_POW5(i) kPow5_128_host[i]
#endif

LH_HD void mul64x64(uint64_t a, uint64_t b, uint64_t* hi, uint64_t* lo) {
#if defined(__CUDA_ARCH__)
  *lo = a * b;
  *hi = __umul64hi(a, b);
#else
  const unsigned __int128 p = (unsigned __int128)a * b;
  *lo = (uint64_t)p;
  *hi = (uint64_t)(p >> 64);
#endif
}
LH_HD int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return __builtin_clzll(x);
#endif
}

// w * 10^q -> the bits of the correctly rounded double; false: not decided here (zero/inf/subnormal result).
// Requires w != 0.
LH_HD bool eisel_lemire(uint64_t w, int q, uint64_t* bits) {
  if (q < -342 || q > 308) return false;
  const int lz = clz64(w);
  w <<= lz;
  const int idx = 2 * (q + 342);
  uint64_t fh, fl;
  mul64x64(w, LH_POW5(idx), &fh, &fl);
  if ((fh & 0x1FFull) == 0x1FFull) {  // the 55 bits that decide the rounding may still change: use the low half
    uint64_t sh, sl;
    mul64x64(w, LH_POW5(idx + 1), &sh, &sl);
    fl += sh;
    if (sh > fl) ++fh;
  }
  if (fl == 0xFFFFFFFFFFFFFFFFull && (q < -27 || q > 55)) return false;  // (conservative; the 2023 proof shows it is not needed)
  const int upperbit = (int)(fh >> 63);
  const int shift = upperbit + 9;  // 64 - 52 - 3
  uint64_t mantissa = fh >> shift;
  int power2 = (int)(((int64_t)217706 * q) >> 16) + 63 + upperbit - lz + 1023;
  if (power2 <= 0) return false;  // subnormal
  // exactly half-way between two doubles: round to even (only possible for small powers of five)
  if (fl <= 1 && q >= -4 && q <= 23 && (mantissa & 3) == 1 && (mantissa << shift) == fh) mantissa &= ~1ull;
  mantissa += mantissa & 1;
  mantissa >>= 1;
  if (mantissa >= (2ull << 52)) {
    mantissa = 1ull << 52;
    ++power2;
  }
  mantissa &= ~(1ull << 52);
  if (power2 >= 0x7FF) return false;  // overflow
  *bits = mantissa | (uint64_t)power2 << 52;
  return true;
}

enum { kDecimalOk = 0, kDecimalNone = 1, kDecimalHard = 2 };

// Reads a number through `next` (a callable returning the current character and `adv` advancing it).
// Generic over the character source so that the device parser can feed it from its register window.
template <class Peek, class Advance>
LH_HD int parse_decimal(Peek peek, Advance adv, double* out) {
  int c = peek();
  bool negative = false;
  if (c == '-') {
    negative = true;
    adv();
    c = peek();
  }
  uint64_t w = 0;
  int digits = 0;       // significant digits taken into w
  int dropped = 0;      // significant digits beyond the 19th
  int exp10 = 0;
  bool any = false;
  while (c >= '0' && c <= '9') {
    any = true;
    if (w != 0 || c != '0') {
      if (digits < 19) { w = w * 10 + (uint64_t)(c - '0'); ++digits; }
      else { ++dropped; ++exp10; }
    }
    adv();
    c = peek();
  }
  if (c == '.') {
    adv();
    c = peek();
    while (c >= '0' && c <= '9') {
      any = true;
      if (w != 0 || c != '0') {
        if (digits < 19) { w = w * 10 + (uint64_t)(c - '0'); ++digits; --exp10; }
        else ++dropped;
      } else {
        --exp10;
      }
      adv();
      c = peek();
    }
  }
  if (!any) return negative ? kDecimalHard : kDecimalNone;
  if (c == 'e' || c == 'E') {
    // an exponent needs at least one digit, else the 'e' is not part of the number (and the caller will
    // find a non-delimiter after it): leave those spellings to the host
    adv();
    c = peek();
    bool eneg = false;
    if (c == '-' || c == '+') {
      eneg = c == '-';
      adv();
      c = peek();
    }
    if (!(c >= '0' && c <= '9')) return kDecimalHard;
    int e = 0;
    while (c >= '0' && c <= '9') {
      if (e < 100000) e = e * 10 + (c - '0');
      adv();
      c = peek();
    }
    exp10 += eneg ? -e : e;
  }
  if (dropped) return kDecimalHard;
  if (w == 0) {
    *out = negative ? -0.0 : 0.0;
    return kDecimalOk;
  }
  uint64_t bits;
  if (!eisel_lemire(w, exp10, &bits)) return kDecimalHard;
  if (negative) bits |= 1ull << 63;
#if defined(__CUDA_ARCH__)
  *out = __longlong_as_double((long long)bits);
#else
  __builtin_memcpy(out, &bits, 8);
#endif
  return kDecimalOk;
}

}  // namespace lhb
