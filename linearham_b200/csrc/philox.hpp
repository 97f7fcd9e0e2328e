// Philox4x32-10 (J. Salmon, M. Moraes, R. Dror, D. Shaw, "Parallel random numbers: as easy as 1, 2, 3", SC'11):
// a counter-based generator - the random words are a pure function of (key, counter), so every (tree, site, draw)
// of the ancestral-state sampler (asr.cu) has its own stream without any state to carry, and the CPU restatement in
// the tests reproduces the device's draws word for word. Pinned to the published known-answer vectors
// (Random123 kat_vectors) in tests/test_asr_oracle.py.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define LH_PHILOX_HD __host__ __device__ __forceinline__
#else
#define LH_PHILOX_HD inline
#endif

namespace lhb {

struct Philox4 {
  uint32_t v[4];
};

LH_PHILOX_HD void philox_mulhilo(uint32_t a, uint32_t b, uint32_t* hi, uint32_t* lo) {
  const uint64_t p = (uint64_t)a * b;
  *hi = (uint32_t)(p >> 32);
  *lo = (uint32_t)p;
}

LH_PHILOX_HD Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int round = 0; round < 10; ++round) {
    uint32_t hi0, lo0, hi1, lo1;
    philox_mulhilo(M0, c0, &hi0, &lo0);
    philox_mulhilo(M1, c2, &hi1, &lo1);
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0;
    c1 = lo1;
    c2 = n2;
    c3 = lo0;
    k0 += W0;
    k1 += W1;
  }
  Philox4 r;
  r.v[0] = c0;
  r.v[1] = c1;
  r.v[2] = c2;
  r.v[3] = c3;
  return r;
}

// a uniform double in (0, 1) from one 32-bit word: (x + 1/2) / 2^32 - exact in double, never 0 or 1
LH_PHILOX_HD double philox_uniform(uint32_t x) { return ((double)x + 0.5) * (1.0 / 4294967296.0); }

}  // namespace lhb
