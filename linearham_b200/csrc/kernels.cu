// Hand-written sm_100a kernels for the phylo-HMM likelihood path.
//
//   k_eigen    GTR eigen system per tree           (replaces libpll models.c:182-331)
//   k_pmatrix  per-branch transition records       (replaces core_pmatrix.c:160-237 and the
//                                                   tip-CLV expansion pll.c:905-964)
//   k_prune    Felsenstein pruning over the xMSA + edge likelihood + naive correction
//              (replaces core_partials.c:575-663, core_likelihood_avx.c:1130-1260,
//               src/PhyloHMM.cpp:228-237)
//   k_hmm      emission fill + scaled forward pass + log-likelihood
//              (replaces src/PhyloHMM.cpp:158-215, src/HMM.cpp:254-319,1107-1177,345-354)
//
// All arithmetic is FP64 on the CUDA cores: the per-site work is 4x4 matrix-vector
// products against per-branch matrices (no dense contraction for tcgen05, which has no
// FP64 kind), and the HMM transition matrices are ~5 % dense.
#include "kernels.hpp"

#include <cstddef>

#include <cfloat>
#include <cmath>

namespace lhb {

#define LH_SCALE_FACTOR 0x1p+256             /* 2^256  (utils.hpp:22, pll.h:89-93) */
#define LH_SCALE_THRESHOLD 0x1p-256         /* 2^-256 */

// ------------------------------------------------------------------ k_eigen ---

// Cyclic Jacobi on the symmetrised rate matrix. The reference uses Householder + QL
// (models.c:24-178); any orthogonal eigensolver gives the same P-matrices to ~1e-16
// because P = I + sum_m inv[.,m] expm1(..) evecs[m,.] is invariant to the choice of
// basis inside a degenerate eigenspace.
__global__ void k_eigen(BatchDev b, int tree0, int ntrees) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntrees) return;
  t += tree0;
  double r[6], pi[4];
  for (int i = 0; i < 6; ++i) r[i] = b.er[6 * t + i];
  for (int i = 0; i < 4; ++i) pi[i] = b.pi[4 * t + i];
  if (r[5] > 0.0) {  // models.c:198-200
    double d = r[5];
    for (int i = 0; i < 6; ++i) r[i] /= d;
  }
  double a[4][4], v[4][4];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) { a[i][j] = 0.0; v[i][j] = (i == j) ? 1.0 : 0.0; }
  int k = 0;
  for (int i = 0; i < 4; ++i)
    for (int j = i + 1; j < 4; ++j) {  // models.c:225-237
      double f = r[k++];
      a[i][j] = a[j][i] = f * sqrt(pi[i] * pi[j]);
      a[i][i] -= f * pi[j];
      a[j][j] -= f * pi[i];
    }
  double mean = 0;
  for (int i = 0; i < 4; ++i) mean += pi[i] * (-a[i][i]);
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) a[i][j] /= mean;  // models.c:240-244

  for (int sweep = 0; sweep < 64; ++sweep) {
    double off = 0;
    for (int p = 0; p < 4; ++p)
      for (int q = p + 1; q < 4; ++q) off += fabs(a[p][q]);
    if (off == 0.0) break;
    for (int p = 0; p < 4; ++p)
      for (int q = p + 1; q < 4; ++q) {
        double apq = a[p][q];
        if (apq == 0.0) continue;
        double g = 100.0 * fabs(apq);
        if (sweep > 4 && fabs(a[p][p]) + g == fabs(a[p][p]) && fabs(a[q][q]) + g == fabs(a[q][q])) {
          a[p][q] = a[q][p] = 0.0;
          continue;
        }
        double h = a[q][q] - a[p][p], tt;
        if (fabs(h) + g == fabs(h)) {
          tt = apq / h;
        } else {
          double theta = 0.5 * h / apq;
          tt = 1.0 / (fabs(theta) + sqrt(1.0 + theta * theta));
          if (theta < 0.0) tt = -tt;
        }
        double c = 1.0 / sqrt(1.0 + tt * tt), s = tt * c, tau = s / (1.0 + c);
        a[p][p] -= tt * apq;
        a[q][q] += tt * apq;
        a[p][q] = a[q][p] = 0.0;
        for (int m = 0; m < 4; ++m) {
          if (m != p && m != q) {
            double amp = a[m][p], amq = a[m][q];
            a[m][p] = a[p][m] = amp - s * (amq + amp * tau);
            a[m][q] = a[q][m] = amq + s * (amp - amq * tau);
          }
          double vmp = v[m][p], vmq = v[m][q];
          v[m][p] = vmp - s * (vmq + vmp * tau);
          v[m][q] = vmq + s * (vmp - vmq * tau);
        }
      }
  }
  double* e = b.eig + (size_t)t * kEigStride;
  for (int m = 0; m < 4; ++m) e[m] = a[m][m];
  for (int m = 0; m < 4; ++m)
    for (int kk = 0; kk < 4; ++kk) {
      e[4 + m * 4 + kk] = v[kk][m] * sqrt(pi[kk]);   // evecs[m][k]   (models.c:317-320)
      e[20 + kk * 4 + m] = v[kk][m] / sqrt(pi[kk]);  // inv_evecs[k][m] (models.c:300-315)
    }
  for (int i = 0; i < 4; ++i) {
    e[36 + i] = log(pi[i]);
    e[40 + i] = pi[i];
  }
}

void launch_eigen(const BatchDev& b, int tree0, int ntrees, cudaStream_t s) {
  int threads = 128;
  k_eigen<<<(ntrees + threads - 1) / threads, threads, 0, s>>>(b, tree0, ntrees);
}

// ---------------------------------------------------------------- k_pmatrix ---
// One thread per (tree, branch, rate category). Branch records are laid out in the order
// the pruning program consumes them, one contiguous stream per (tree, category): record
// 2*step+c for child c of a step, then the root edge, then an identity matrix (the edge
// evaluation runs as one more step: the root CLV "crosses" the identity); a record is 20 doubles:
//   inner child : P[i][j] at i*4 + j                (row-major as core_pmatrix.c; 4 pad)
//   tip child   : Q[code][i] at code*4 + i, code in A,C,G,T,N: the product
//                 P[k] * tipclv(code) that set_tipclv + the generic kernel would form
//                 (pll.c:926-930: one-hot for ACGT, all ones for N).
constexpr int kRecDoublesPm = 20;  // doubles per branch record (kRecDoubles)

__global__ void __launch_bounds__(128) k_pmatrix(FamilyDev f, BatchDev b, int tree0, int ntrees) {
  // A warp takes 32 consecutive branches of one (tree, category) pass: their records are 5120
  // contiguous bytes of the stream, written with full 512-byte warp stores after a transpose
  // through shared memory (a thread storing its own 160-byte record would touch ten half-used sectors).
  __shared__ double tile[4][32 * 21];  // per warp: 32 records x 20 doubles, pitch 21 (conflict-free)
  const int NB = 2 * f.T - 3, NR = records_per_pass(f.T), R = b.R;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int groups = (NR + 31) / 32;  // record groups per pass
  const long long total = (long long)ntrees * R * groups;
  double* my = tile[wid];
  for (long long g = (long long)blockIdx.x * 4 + wid; g < total; g += (long long)gridDim.x * 4) {
    const int bg = (int)(g % groups);
    const int k = (int)((g / groups) % R);
    const int lt = (int)(g / ((long long)groups * R));
    const int t = tree0 + lt;
    const int br = bg * 32 + lane;
    __syncwarp();
    if (br < NR) {
      const int step = min(br, NB - 1) >> 1;
      const int4 st = __ldg(b.plan + (size_t)t * (f.T - 1) + step);
      const int src = (br & 1) ? st.y : st.x;  // for the final step y is the child side
      const bool tip = br < NB && ((br == NB - 1) ? (st.y >= 0) : (src >= 0));
      const double* e = b.eig + (size_t)t * kEigStride;
      const double len = br < NB ? b.pblen[(size_t)t * NB + br] : 0.0;
      const double rate = b.rates[(size_t)t * R + k];
      double P[4][4];
      if (len == 0.0) {  // core_pmatrix.c:174-179 (and the closing identity record)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int c = 0; c < 4; ++c) P[j][c] = (j == c) ? 1.0 : 0.0;
      } else {
        double expd[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) expd[m] = expm1(e[m] * rate * len);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          double tmp[4];
#pragma unroll
          for (int m = 0; m < 4; ++m) tmp[m] = e[20 + j * 4 + m] * expd[m];
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            double acc = (j == c) ? 1.0 : 0.0;
#pragma unroll
            for (int m = 0; m < 4; ++m) acc += tmp[m] * e[4 + m * 4 + c];
            P[j][c] = acc;
          }
        }
      }
      double* o = my + lane * 21;
      if (!tip) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int c = 0; c < 4; ++c) o[j * 4 + c] = P[j][c];
#pragma unroll
        for (int c = 0; c < 4; ++c) o[16 + c] = 0.0;
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int j = 0; j < 4; ++j) o[c * 4 + j] = P[j][c];
#pragma unroll
        for (int i = 0; i < 4; ++i) o[16 + i] = ((P[i][0] + P[i][1]) + P[i][2]) + P[i][3];
      }
    }
    __syncwarp();
    // the group's records are contiguous: doubles [first, first + n*20) of the pass stream
    const int n_rec = min(32, NR - bg * 32);
    double* dst = b.prec + (((size_t)lt * R + k) * NR + (size_t)bg * 32) * kRecDoublesPm;
    for (int i = lane; i < n_rec * kRecDoublesPm; i += 32) dst[i] = my[(i / kRecDoublesPm) * 21 + (i % kRecDoublesPm)];
  }
}

void launch_pmatrix(const FamilyDev& f, const BatchDev& b, int tree0, int ntrees, cudaStream_t s) {
  const int groups = (records_per_pass(f.T) + 31) / 32;
  long long total = (long long)ntrees * b.R * groups;  // warps of work
  long long blocks = (total + 3) / 4;
  if (blocks > 148LL * 64) blocks = 148LL * 64;
  k_pmatrix<<<(unsigned)blocks, 128, 0, s>>>(f, b, tree0, ntrees);
}

// ------------------------------------------------------------------ k_prune ---
// One warp per CTA, persistent; a work item is (tree, tile of 32*NS xMSA columns); lane l owns
// columns tile_base + u*32 + l (u < NS) and walks the tree's program (plan.hpp) once per rate
// category ("category-outer"): the categories of a site never mix before the edge likelihood,
// so a pass needs only 4 doubles of state per column. That lets one thread carry NS columns in
// registers, and the 4x4 transition matrix of a branch, read once into registers, serves NS
// columns. (With one column per thread the warp-uniform matrix loads saturated the L1 data pipe,
// not the FP64 pipe: ncu r1a, lsu wavefronts 57 %, fp64 14 %.)
//
// Nothing in a step waits on L2 (ncu r1a: 13.9 warps stalled on long scoreboard per issue):
//   plan ring   64 steps in shared memory, refilled 32 at a time with cp.async
//   record ring 4 steps x 2 branches x 20 doubles (the category's P or Q blocks; the records of a
//               pass are one contiguous stream, 320 bytes per step), cp.async two steps ahead
//   tip codes   one 32-bit load per tip child (the NS codes of the lane, pre-multiplied by 32 =
//               byte offset of the Q column), issued one step ahead into registers
//
// Inner CLVs never exist as arrays in HBM: the running result stays in registers, results
// parked for a later step go to a per-column stack of which `smem_slots` levels (the busiest) are shared
// memory ([slot][u][state][lane], conflict-free because a column is only ever touched by its own
// thread) and whose deeper levels are an L2-resident scratch ring private to the CTA.
//
// Scaling. libpll rescales a site by 2^256 when ALL 4R entries of a node are below 2^-256
// (core_partials.c:648-659). Here each category keeps its own count and rescales when its own 4
// entries are below the threshold; the counts are reconciled exactly (powers of two) when the
// categories are summed at the root edge. The two schemes give the same doubles whenever the
// reference's unscaled categories stay normal numbers; where they would go subnormal in the
// reference, this kernel keeps their full precision (a relative contribution < 2^-256).

#ifndef LH_PRUNE_MIN_CTAS
#define LH_PRUNE_MIN_CTAS 12
#define LH_PRUNE_MIN_CTAS3 12
#define LH_PRUNE_MIN_CTAS2 16
#define LH_PRUNE_MIN_CTAS1 24
#endif

constexpr int kBlk = 4;        // k_prune: steps per staged block of branch records (two blocks resident)
constexpr int kPlanRing = 64;  // plan entries resident
constexpr int kRecDoubles = 20;
constexpr int kStepBytes = 2 * kRecDoubles * 8;  // both branches of a step, one category

__device__ __forceinline__ void cp_async16(unsigned smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_dst), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// Bulk asynchronous copies (one instruction moves a whole staged block; completion is counted in
// bytes on an mbarrier that every lane then waits on).
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "LH_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra LH_DONE;\n"
      "bra LH_WAIT;\n"
      "LH_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}

__device__ __forceinline__ unsigned shared_address_opaque(const void* p) {
  unsigned a;
  asm volatile("{\n.reg .u64 t;\ncvta.to.shared.u64 t, %1;\ncvt.u32.u64 %0, t;\n}\n" : "=r"(a) : "l"(p));
  return a;
}

// Shared-memory accesses by 32-bit shared-space address: through generic pointers every access
// site first rebuilds the CTA's shared window base (S2UR SR_CgaCtaId, UMOV, ULEA).
__device__ __forceinline__ double2 lds_f64x2(unsigned a) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ double lds_f64(unsigned a) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ int lds_s32(unsigned a) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ int4 lds_s32x4(unsigned a) {
  int4 v;
  asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f64(unsigned a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v)); }
__device__ __forceinline__ void sts_s32(unsigned a, int v) { asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v)); }

// v < 2^-256 on the high word: 2^-256 = 0x2FF0000000000000, so for v >= +0 the comparison of the
// bit patterns reduces to the high words, and every negative v (sign bit set) is below the
// threshold as a signed integer too. Keeps the test off the FP64 pipe.
constexpr int kThresholdHi = 0x2FF00000;

// Parked CLV levels of this lane: [slot][u][state] doubles at stride 32 and [slot][u] counts at
// stride 32; the levels smem_lo .. smem_lo + smem_slots - 1 live in shared memory (addressed by
// 32-bit shared-space address), the others in the CTA's global scratch.
template <int NS>
struct SlotIO {
  unsigned s_clv, s_cnt;  // shared-space byte addresses of this lane's element of level smem_lo
  double* g_clv;          // [gslots][NS*4][32] (this CTA), already offset by the lane
  int* g_cnt;             // [gslots][NS][32]
  int smem_slots, smem_lo;

  __device__ __forceinline__ bool in_smem(int slot) const { return (unsigned)(slot - smem_lo) < (unsigned)smem_slots; }
  __device__ __forceinline__ unsigned shared_clv(int slot) const { return s_clv + (unsigned)(slot - smem_lo) * (4 * NS * 32 * 8); }
  __device__ __forceinline__ unsigned shared_cnt(int slot) const { return s_cnt + (unsigned)(slot - smem_lo) * (NS * 32 * 4); }
  __device__ __forceinline__ double* global_clv(int slot) const {
    return g_clv + (size_t)(slot < smem_lo ? slot : slot - smem_slots) * (4 * NS) * 32;
  }
  __device__ __forceinline__ int* global_cnt(int slot) const {
    return g_cnt + (slot < smem_lo ? slot : slot - smem_slots) * NS * 32;
  }
};

// m[u][i] = sum_j P[i][j] * c[u][j]   (core_partials.c:617-624), P row-major in shared memory
__device__ __forceinline__ void load_p(unsigned rec, double (&P)[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const double2 p01 = lds_f64x2(rec + 32 * i);
    const double2 p23 = lds_f64x2(rec + 32 * i + 16);
    P[i][0] = p01.x; P[i][1] = p01.y; P[i][2] = p23.x; P[i][3] = p23.y;
  }
}
__device__ __forceinline__ void matvec(const double (&P)[4][4], const double (&c)[4], double (&m)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    double acc = P[i][0] * c[0];
    acc += P[i][1] * c[1];
    acc += P[i][2] * c[2];
    acc += P[i][3] * c[3];
    m[i] = acc;
  }
}

// Messages of a tip child for the NS columns of this thread: column `code` of the staged Q block
// (`codes` holds the lane's NS byte offsets into it; `rec` is the block's shared-space address).
template <int NS>
__device__ __forceinline__ void tip_messages(unsigned rec, unsigned codes, double (&m)[NS][4]) {
#pragma unroll
  for (int u = 0; u < NS; ++u) {
    const unsigned off = (codes >> (8 * u)) & 0xffu;
    const double2 q01 = lds_f64x2(rec + off);
    const double2 q23 = lds_f64x2(rec + off + 16);
    m[u][0] = q01.x; m[u][1] = q01.y; m[u][2] = q23.x; m[u][3] = q23.y;
  }
}

template <int NS>
__device__ __forceinline__ void park(const SlotIO<NS>& io, int slot, const double (&reg)[NS][4], const int (&reg_cnt)[NS]) {
  if (io.in_smem(slot)) {
    const unsigned clv = io.shared_clv(slot), sc = io.shared_cnt(slot);
#pragma unroll
    for (int u = 0; u < NS; ++u) {
#pragma unroll
      for (int i = 0; i < 4; ++i) sts_f64(clv + (u * 4 + i) * 32 * 8, reg[u][i]);
      sts_s32(sc + u * 32 * 4, reg_cnt[u]);
    }
  } else {
    double* clv = io.global_clv(slot);
    int* sc = io.global_cnt(slot);
#pragma unroll
    for (int u = 0; u < NS; ++u) {
#pragma unroll
      for (int i = 0; i < 4; ++i) clv[(u * 4 + i) * 32] = reg[u][i];
      sc[u * 32] = reg_cnt[u];
    }
  }
}
// dynamic part: the parked CLV levels; static part: the two rings (compile-time addresses, so the
// ring reads need no base register)
static size_t prune_dynamic_smem_bytes(int ns, int smem_slots) {
  return (size_t)smem_slots * 32 * ns * (4 * sizeof(double) + sizeof(int));
}
size_t prune_smem_bytes(int ns, int smem_slots) {
  return (size_t)2 * kBlk * kStepBytes + kPlanRing * sizeof(int4) + 4 * sizeof(unsigned long long) +
         prune_dynamic_smem_bytes(ns, smem_slots);
}

template <int NS>
__global__ void __launch_bounds__(32, (NS == 4 ? LH_PRUNE_MIN_CTAS : NS == 3 ? LH_PRUNE_MIN_CTAS3 : NS == 2 ? LH_PRUNE_MIN_CTAS2 : LH_PRUNE_MIN_CTAS1))
    k_prune(FamilyDev f, BatchDev b, PruneCfg cfg, int tree0, int ntrees) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x, T = f.T, R = b.R;
  const int NB = 2 * T - 3, n_steps = T - 2, n_iter = T - 1;  // node updates + the edge evaluation
  const int gslots = cfg.total_slots - cfg.smem_slots;
  struct __align__(16) Rings {
    double rec[2 * kBlk * 2 * kRecDoubles];  // two blocks of kBlk steps of branch records
    int4 plan[kPlanRing];                    // two halves of 32 steps
    unsigned long long bars[4];              // mbarriers: records 0/1, plan 0/1
  };
  __shared__ Rings rings;
  // One register holds the shared-space address everything static hangs off, one the dynamic part's.
  // (Opaque to the compiler on purpose: it would otherwise rebuild the window base at every use.)
  const unsigned rings_s = shared_address_opaque(&rings);
  const unsigned ring_rec_0 = rings_s + (unsigned)offsetof(Rings, rec);
  const unsigned ring_plan_s = rings_s + (unsigned)offsetof(Rings, plan);
  const unsigned bars_s = rings_s + (unsigned)offsetof(Rings, bars);
  SlotIO<NS> io;
  io.s_clv = shared_address_opaque(smem_raw) + lane * 8;
  io.s_cnt = io.s_clv + (unsigned)cfg.smem_slots * (4 * NS * 32 * 8) - lane * 4;
  io.g_clv = cfg.scratch + (size_t)blockIdx.x * (gslots > 0 ? gslots : 0) * (4 * NS) * 32 + lane;
  io.g_cnt = cfg.scratch_cnt + (size_t)blockIdx.x * (gslots > 0 ? gslots : 0) * NS * 32 + lane;
  io.smem_slots = cfg.smem_slots;
  io.smem_lo = cfg.smem_lo;
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) mbar_init(bars_s + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  unsigned phase = 0;  // bit i: the parity the next wait on bars[i] looks for
  constexpr unsigned kBlkBytes = kBlk * kStepBytes, kHalfBytes = 32 * sizeof(int4);
  const int n_blocks = (n_iter + kBlk - 1) / kBlk, n_halves = (n_iter + 31) / 32;
  const size_t pass_bytes = (size_t)records_per_pass(T) * kRecDoubles * 8;  // records of one (tree, category)
  const unsigned code_pitch = (unsigned)cfg.n_tiles * 32u;  // words per tip row (the table is far below 2^32 words)

  // work item = (tree, tile, category): the categories of a tile are independent until k_edge sums
  // them, and items this fine keep the last wave of a small batch full
  const int per_tree = cfg.n_tiles * R;
  // text input: the chunk's list holds the pattern trees first; how many was decided on the device
  const int32_t* tree_list = cfg.tree_list;
  if (cfg.dyn_n_pat) {
    const int n_pat = __ldg(cfg.dyn_n_pat);
    if (cfg.defer_naive) {
      ntrees = n_pat;
    } else {
      ntrees = cfg.dyn_total - n_pat;
      tree_list += n_pat;
    }
  }
  const int n_work = ntrees * per_tree;
  for (int w = blockIdx.x; w < n_work;) {
    // claim the next item now; the answer is needed when this one is done
    int w_next = 0;
    if (lane == 0) w_next = (int)gridDim.x + atomicAdd(cfg.work_counter, 1);
    const int li = w / per_tree, rem = w - li * per_tree;
    const int tile = rem / R, k = rem - tile * R;
    const int t = __ldg(tree_list + li);
    const int lt = t - tree0;
    const int4* __restrict__ plan = b.plan + (size_t)t * (T - 1);
    const unsigned* __restrict__ my_codes = cfg.codes_t + (size_t)tile * 32 + lane;
    int naive_br = -1;  // record index of the naive tip's branch when its factor is deferred

    {
      // The record stream of this (tree, category) - step q at rec_src + q * 320 bytes, the last
      // step a single record - is staged kBlk steps per bulk copy, two blocks resident; the plan
      // 32 steps per copy, two halves resident. Lane 0 issues, every lane waits.
      const char* rec_src = reinterpret_cast<const char*>(b.prec) + ((size_t)lt * R + k) * pass_bytes;
      auto issue_records = [&](int jb) {
        const unsigned off = (unsigned)jb * kBlkBytes;
        const unsigned bytes = min(kBlkBytes, (unsigned)pass_bytes - off);
        const unsigned bar = bars_s + 8 * (jb & 1);
        mbar_expect_tx(bar, bytes);
        bulk_g2s(ring_rec_0 + (jb & 1) * kBlkBytes, rec_src + off, bytes, bar);
      };
      auto issue_plan = [&](int h) {
        const unsigned bytes = (unsigned)min(32, n_iter - 32 * h) * (unsigned)sizeof(int4);
        const unsigned bar = bars_s + 16 + 8 * (h & 1);
        mbar_expect_tx(bar, bytes);
        bulk_g2s(ring_plan_s + (h & 1) * kHalfBytes, plan + 32 * h, bytes, bar);
      };
      auto wait_on = [&](int i) {
        mbar_wait(bars_s + 8 * i, (phase >> i) & 1u);
        phase ^= 1u << i;
      };
      __syncwarp();  // the previous work item is done with the rings
      if (lane == 0) {
        issue_plan(0);
        if (n_halves > 1) issue_plan(1);
        issue_records(0);
        if (n_blocks > 1) issue_records(1);
      }
      wait_on(2);

      double reg[NS][4];
      int reg_cnt[NS];
#pragma unroll
      for (int u = 0; u < NS; ++u) {
        reg_cnt[u] = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) reg[u][i] = 0.0;
      }
      // software pipeline registers: the step being executed and the next one
      int4 st_next = lds_s32x4(ring_plan_s);
      unsigned cA_next = 0, cB_next = 0;
      if (st_next.x >= 0) cA_next = __ldg(my_codes + (unsigned)st_next.x * code_pitch);
      if (st_next.y >= 0) cB_next = __ldg(my_codes + (unsigned)st_next.y * code_pitch);

      for (int jb = 0; jb < n_blocks; ++jb) {
        wait_on(jb & 1);  // the records of this block
        // the step before a new plan half reads that half's first entry
        if ((jb & 7) == 7 && (jb >> 3) + 1 < n_halves) wait_on(2 + (((jb >> 3) + 1) & 1));
        const int s_end = min(n_iter, (jb + 1) * kBlk);
        unsigned recA = ring_rec_0 + (jb & 1) * kBlkBytes;
#pragma unroll 1
        for (int s = jb * kBlk; s < s_end; ++s, recA += kStepBytes) {
        const int4 st = st_next;
        const unsigned cA = cA_next, cB = cB_next;
        if (s + 1 < n_iter) {     // plan and tip codes of the next step
          st_next = lds_s32x4(ring_plan_s + ((s + 1) & (kPlanRing - 1)) * 16);
          if (st_next.x >= 0) cA_next = __ldg(my_codes + (unsigned)st_next.x * code_pitch);
          if (st_next.y >= 0) cB_next = __ldg(my_codes + (unsigned)st_next.y * code_pitch);
        }
        // A step combines two messages (core_partials.c:617-628). At most one child is the running
        // result (REG) and a parked child only ever meets REG, so a step is: the message of the
        // "other" child (a tip's column lookup, or a parked CLV through its matrix), times REG through
        // its matrix - or a second lookup for a cherry. The edge evaluation is one more step of the
        // same shape: its records are the root edge (child side, st.y) and an identity (root side,
        // st.x), and the two messages go to k_edge instead of being multiplied.
        const bool last = s == n_steps;
        const unsigned recX = last ? recA + kRecDoubles * 8 : recA;  // record of child st.x
        const unsigned recY = last ? recA : recA + kRecDoubles * 8;  // record of child st.y
        const bool a_reg = st.x == -1, any_reg = a_reg | (st.y == -1);
        const int o_src = a_reg ? st.y : st.x;
        const unsigned o_rec = a_reg ? recY : recX, o_codes = a_reg ? cB : cA;
        double m[NS][4];
        // Pattern columns: the node under which the naive tip (row 0) hangs keeps the other child's
        // message only (a factor of exactly 1 stands in for the tip's); k_edge multiplies the naive
        // tip's Q column in, per xMSA column, and applies the node's rescaling test then.
        bool deferred = false;
        if (o_src >= 0) {
          if (cfg.defer_naive && o_src == 0) {
            deferred = true;
            naive_br = 2 * s + (o_rec != recA ? 1 : 0);
#pragma unroll
            for (int u = 0; u < NS; ++u)
#pragma unroll
              for (int i = 0; i < 4; ++i) m[u][i] = 1.0;
          } else {
            tip_messages<NS>(o_rec, o_codes, m);
          }
        } else {
          // a parked CLV through its matrix; its scaler counts join REG's right away (a parked child
          // only ever meets REG; fill_parent_scaler, core_partials.c:24-46)
          double P[4][4];
          load_p(o_rec, P);
          const int slot = -2 - o_src;
          if (io.in_smem(slot)) {
            const unsigned clv = io.shared_clv(slot), sc = io.shared_cnt(slot);
#pragma unroll
            for (int u = 0; u < NS; ++u) {
              double c[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) c[j] = lds_f64(clv + (u * 4 + j) * 32 * 8);
              matvec(P, c, m[u]);
              reg_cnt[u] += lds_s32(sc + u * 32 * 4);
            }
          } else {
            const double* clv = io.global_clv(slot);
            const int* sc = io.global_cnt(slot);
#pragma unroll
            for (int u = 0; u < NS; ++u) {
              double c[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) c[j] = clv[(u * 4 + j) * 32];
              matvec(P, c, m[u]);
              reg_cnt[u] += sc[u * 32];
            }
          }
        }
        if (any_reg) {
          double P[4][4];
          load_p(a_reg ? recX : recY, P);
          if (!last) {
#pragma unroll
            for (int u = 0; u < NS; ++u) {
              double mr[4];
              matvec(P, reg[u], mr);
#pragma unroll
              for (int i = 0; i < 4; ++i) reg[u][i] = m[u][i] * mr[i];  // core_partials.c:626
            }
          } else {
            // ---- edge term of this category (core_likelihood_avx.c:1130-1239): hand its inputs to
            // k_edge, rows 0..3 the message of the node across the edge, rows 4..7 the root CLV ----
            double* eo = cfg.ebuf + (size_t)lt * cfg.e_stride + (size_t)k * 8 * cfg.cols_pad + tile * (32 * NS) + lane;
            int* co = cfg.ecnt + (size_t)lt * cfg.c_stride + (size_t)k * cfg.cols_pad + tile * (32 * NS) + lane;
            double* e_other = eo + (a_reg ? (size_t)0 : (size_t)4 * cfg.cols_pad);  // st.x is the root side
            double* e_reg = eo + (a_reg ? (size_t)4 * cfg.cols_pad : (size_t)0);
#pragma unroll
            for (int u = 0; u < NS; ++u) {
              double mr[4];
              matvec(P, reg[u], mr);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                e_other[(size_t)i * cfg.cols_pad + u * 32] = m[u][i];
                e_reg[(size_t)i * cfg.cols_pad + u * 32] = mr[i];
              }
              co[u * 32] = reg_cnt[u];
            }
          }
        } else {  // a cherry: two tips
          double m2[NS][4];
          if (cfg.defer_naive && st.y == 0) {
            deferred = true;
            naive_br = 2 * s + 1;
#pragma unroll
            for (int u = 0; u < NS; ++u)
#pragma unroll
              for (int i = 0; i < 4; ++i) m2[u][i] = 1.0;
          } else {
            tip_messages<NS>(recY, cB, m2);
          }
#pragma unroll
          for (int u = 0; u < NS; ++u) {
#pragma unroll
            for (int i = 0; i < 4; ++i) reg[u][i] = m[u][i] * m2[u][i];
            reg_cnt[u] = 0;
          }
        }
        if (!last) {
          int lowest = 0x7fffffff;  // min over columns of the largest high word
#pragma unroll
          for (int u = 0; u < NS; ++u) {
            const int hi = max(max(__double2hiint(reg[u][0]), __double2hiint(reg[u][1])),
                               max(__double2hiint(reg[u][2]), __double2hiint(reg[u][3])));
            lowest = min(lowest, hi);
          }
          if (lowest < kThresholdHi && !deferred) {  // some column has all four entries below 2^-256
#pragma unroll
            for (int u = 0; u < NS; ++u) {
              const int hi = max(max(__double2hiint(reg[u][0]), __double2hiint(reg[u][1])),
                                 max(__double2hiint(reg[u][2]), __double2hiint(reg[u][3])));
              if (hi < kThresholdHi) {
#pragma unroll
                for (int i = 0; i < 4; ++i) reg[u][i] *= LH_SCALE_FACTOR;
                reg_cnt[u] += 1;
              }
            }
          }
          if (st.z >= 0) park<NS>(io, st.z, reg, reg_cnt);
        }
        }
        __syncwarp();  // every lane is done with this block's records (and, at a half's end, its plan entries)
        if (lane == 0) {
          if (jb + 2 < n_blocks) issue_records(jb + 2);
          if ((jb & 7) == 7 && (jb >> 3) + 2 < n_halves) issue_plan((jb >> 3) + 2);
        }
      }
    }

    if (tile == 0 && k == 0 && lane == 0) cfg.enaive[lt] = naive_br;
    w = __shfl_sync(0xffffffffu, w_next, 0);
  }
}

#define LH_DISPATCH_NS(NS_, CALL)                     \
  switch (NS_) {                                      \
    case 1: { constexpr int NN = 1; CALL; } break;    \
    case 2: { constexpr int NN = 2; CALL; } break;    \
    case 3: { constexpr int NN = 3; CALL; } break;    \
    default: { constexpr int NN = 4; CALL; } break;   \
  }

template <int NS>
static void launch_prune_n(const FamilyDev& f, const BatchDev& b, const PruneCfg& c, int tree0,
                           int ntrees, cudaStream_t s) {
  size_t smem = prune_dynamic_smem_bytes(NS, c.smem_slots);
  ensure_dynamic_smem(reinterpret_cast<const void*>(&k_prune<NS>), smem);
  cudaMemsetAsync(c.work_counter, 0, sizeof(int32_t), s);
  k_prune<NS><<<c.grid, 32, smem, s>>>(f, b, c, tree0, ntrees);
}

template <int NS>
static int prune_occ_n(size_t smem) {
  int n = 0;
  ensure_dynamic_smem(reinterpret_cast<const void*>(&k_prune<NS>), smem);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_prune<NS>, 32, smem);
  return n;
}

void launch_prune(const FamilyDev& f, const BatchDev& b, const PruneCfg& c, int tree0, int ntrees,
                  cudaStream_t s) {
  LH_DISPATCH_NS(c.sites_per_thread, launch_prune_n<NN>(f, b, c, tree0, ntrees, s));
}

int prune_max_ctas_per_sm(int ns, int smem_slots) {
  size_t smem = prune_dynamic_smem_bytes(ns, smem_slots);  // (the occupancy query adds the static part)
  int n = 0;
  LH_DISPATCH_NS(ns, n = prune_occ_n<NN>(smem));
  return n;
}

// ------------------------------------------------------------------- k_edge ---
// One thread per (tree, xMSA column): the edge log-likelihood of the column from the two CLV
// messages k_prune left per category (core_likelihood_avx.c:1130-1251), the naive-sequence
// correction and the exponential (src/PhyloHMM.cpp:219-240). When the pruning ran on site
// patterns, the naive tip's factor - column `code` of its branch's Q block - is multiplied in
// here, with the node's rescaling test applied to the product - per rate category (this kernel's rule;
// libpll tests all 4R entries of the site at once, core_partials.c:648-659), the counts being reconciled
// exactly when the categories are summed. tests/test_scaling_regime.py pins the results in the underflow
// regime to the reference's.
__global__ void __launch_bounds__(128) k_edge(FamilyDev f, BatchDev b, EdgeCfg cfg, int tree0, int ntrees) {
  const int S = f.S, R = b.R, NB = 2 * f.T - 3, edge_br = NB - 1;
  const double weight = 1.0 / R;  // pll.c:746-747
  constexpr double kDown = LH_SCALE_THRESHOLD;
  const long long total = (long long)ntrees * S;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int lt = (int)(idx / S), site = (int)(idx - (long long)lt * S);
    const int t = tree0 + lt;
    const int code = __ldg(f.tip_codes + site);  // xMSA row 0 is the naive sequence
    const int nbr = __ldg(cfg.enaive + lt);
    int col = nbr >= 0 ? __ldg(cfg.col_of_site + site) : site;
    const size_t pitch = nbr >= 0 ? cfg.cols_pad_pat : cfg.cols_pad_all;
    const double* eig = b.eig + (size_t)t * kEigStride;
    const double pi0 = eig[40], pi1 = eig[41], pi2 = eig[42], pi3 = eig[43];
    // sum over categories of the edge terms, kept as acc * 2^(-256 * acc_cnt)
    double acc = 0.0;
    int acc_cnt = 0;
    for (int k = 0; k < R; ++k) {
      const double* e = cfg.ebuf + (size_t)lt * cfg.e_stride + (size_t)k * 8 * pitch + col;
      double mc[4], rg[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        mc[i] = e[(size_t)i * pitch];
        rg[i] = e[(size_t)(4 + i) * pitch];
      }
      int c = cfg.ecnt[(size_t)lt * cfg.c_stride + (size_t)k * pitch + col];
      if (nbr >= 0) {
        const double* q = b.prec + (((size_t)lt * R + k) * records_per_pass(f.T) + nbr) * kRecDoubles + code * 4;
        if (nbr == edge_br) {  // the naive tip is the node across the root edge
#pragma unroll
          for (int i = 0; i < 4; ++i) mc[i] = q[i];
        } else {               // the naive tip hangs under the virtual root
#pragma unroll
          for (int i = 0; i < 4; ++i) rg[i] *= q[i];  // core_partials.c:626
          const int hi = max(max(__double2hiint(rg[0]), __double2hiint(rg[1])),
                             max(__double2hiint(rg[2]), __double2hiint(rg[3])));
          if (hi < kThresholdHi) {
#pragma unroll
            for (int i = 0; i < 4; ++i) rg[i] *= LH_SCALE_FACTOR;
            c += 1;
          }
        }
      }
      const double t0 = (pi0 * mc[0]) * rg[0];
      const double t1 = (pi1 * mc[1]) * rg[1];
      const double t2 = (pi2 * mc[2]) * rg[2];
      const double t3 = (pi3 * mc[3]) * rg[3];
      double term = (t0 + t1) + (t2 + t3);
      if (cfg.rate_ll) {
        double v = log(term);  // -inf for 0
        if (c) v += c * cfg.log_threshold;
        if (code != 4) v -= eig[36 + code];
        cfg.rate_ll[((size_t)t * R + k) * S + site] = term > 0.0 ? v : -INFINITY;
      }
      if (term > 0.0) {  // :1225
        term *= weight;
        // bring acc and term to the smaller count (exact: powers of two)
        if (acc == 0.0) acc_cnt = c;
        while (c > acc_cnt) { term *= kDown; --c; if (term == 0.0) c = acc_cnt; }
        while (acc_cnt > c) { acc *= kDown; --acc_cnt; if (acc == 0.0) acc_cnt = c; }
        acc += term;
      }
    }
    double site_lk = log(acc);
    if (acc_cnt) site_lk += acc_cnt * cfg.log_threshold;   // :1247-1251
    // naive-sequence correction (src/PhyloHMM.cpp:229-237)
    if (code != 4) site_lk -= eig[36 + code];
    // xmsa_emission_.array().exp() (src/PhyloHMM.cpp:239) is Eigen's packet exp (pexp_double; SSE2 Packet2d in the
    // reference's build): 2^n * exp(g) with n = floor(x log2(e) + 1/2), which is 0 once n <= -1023 - where libm
    // returns a subnormal; the odd last element of the vector goes through libm
    const bool flush = site < (S & ~1) && floor(fma(site_lk, 1.4426950408889634073599, 0.5)) <= -1023.0;
    b.xe[(size_t)t * S + site] = flush ? 0.0 : exp(site_lk);
  }
}

void launch_edge(const FamilyDev& f, const BatchDev& b, const EdgeCfg& c, int tree0, int ntrees,
                 cudaStream_t s) {
  const long long total = (long long)ntrees * f.S;
  const int threads = 128;
  long long blocks = (total + threads - 1) / threads;
  if (blocks > 148LL * 64) blocks = 148LL * 64;
  k_edge<<<(unsigned)blocks, threads, 0, s>>>(f, b, c, tree0, ntrees);
}

// -------------------------------------------------------------------- k_hmm ---
// One warp per tree sample. Per-warp shared memory: two state vectors (ping-pong) and
// the five germline/padding emission vectors with their scaler counts.

__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int warp_max_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ScaleMatrix (src/utils.cpp:135-144) on a row of n doubles owned by the warp
// (entry j belongs to lane j % 32). Multiplying by 2^256 is exact, so the number of
// rounds the reference's while-loop makes is determined by the smallest positive entry.
// `tame` comes back true when every entry of the scaled row is a number no larger than 1e150.
__device__ __forceinline__ int warp_scale_row(double* row, int n, int lane, bool& tame) {
  double mn = DBL_MAX, mx = 0.0;
  bool odd = false;  // NaN, infinity or a negative entry
  for (int j = lane; j < n; j += 32) {
    const double v = row[j];
    if (v > 0.0 && v < mn) mn = v;
    mx = fmax(mx, v);
    odd |= !(v >= 0.0 && v <= DBL_MAX);
  }
  mn = warp_min(mn);
  mx = -warp_min(-mx);
  int rounds = 0;
  while (mn < LH_SCALE_THRESHOLD) {
    mn *= LH_SCALE_FACTOR;
    mx *= LH_SCALE_FACTOR;
    ++rounds;
  }
  if (rounds) {
    for (int j = lane; j < n; j += 32) {
      double v = row[j];
      for (int r = 0; r < rounds; ++r) v *= LH_SCALE_FACTOR;
      row[j] = v;
    }
  }
  tame = !__any_sync(0xffffffffu, odd) && mx <= 1e150;
  __syncwarp();
  return rounds;
}
__device__ __forceinline__ int warp_scale_row(double* row, int n, int lane) {
  bool tame;
  return warp_scale_row(row, n, lane, tame);
}

// FillGermlinePaddingEmission (src/PhyloHMM.cpp:158-193) for one region: lane g owns
// gene g's running product; returns the region's max scaler count.
__device__ __forceinline__ int fill_region(int n, const int32_t* __restrict__ off,
                                           const int32_t* __restrict__ inds,
                                           const double* __restrict__ xe, double* emis, int lane) {
  int local_max = 0;
  for (int g = lane; g < n; g += 32) {
    double x = 1.0;
    int cnt = 0;
    const int e = off[g + 1];
    for (int j = off[g]; j < e; ++j) {
      x *= xe[inds[j]];
      while (x > 0.0 && x < LH_SCALE_THRESHOLD) {
        x *= LH_SCALE_FACTOR;
        ++cnt;
      }
    }
    emis[g] = x;
    reinterpret_cast<int*>(emis + n)[g] = cnt;  // counts parked behind the values
    local_max = max(local_max, cnt);
  }
  const int mx = warp_max_i(local_max);
  __syncwarp();
  for (int g = lane; g < n; g += 32) {
    const int d = mx - reinterpret_cast<int*>(emis + n)[g];
    // std::pow(SCALE_FACTOR, d): 2^(256 d), +inf from d = 4 on (src/PhyloHMM.cpp:191)
    double pw = 1.0;
    for (int r = 0; r < d; ++r) pw *= LH_SCALE_FACTOR;
    emis[g] *= pw;
  }
  __syncwarp();
  return mx;
}

// row-vector times sparse matrix: out[j] = sum_k prev[k] * M[k][j]
__device__ __forceinline__ double csc_dot(const CscDev& m, const double* prev, int j) {
  // The terms are added in row order (the reference's dense dot product with the zero terms dropped);
  // four entries and their operands are fetched ahead of the four dependent additions.
  double acc = 0.0;
  const int e = m.colptr[j + 1];
  int p = m.colptr[j];
  for (; p + 4 <= e; p += 4) {
    const double2 e0 = __ldg(m.ent + p), e1 = __ldg(m.ent + p + 1), e2 = __ldg(m.ent + p + 2), e3 = __ldg(m.ent + p + 3);
    const double x0 = prev[__double2loint(e0.y)], x1 = prev[__double2loint(e1.y)];
    const double x2 = prev[__double2loint(e2.y)], x3 = prev[__double2loint(e3.y)];
    acc += x0 * e0.x;
    acc += x1 * e1.x;
    acc += x2 * e2.x;
    acc += x3 * e3.x;
  }
  for (; p < e; ++p) {
    const double2 en = __ldg(m.ent + p);
    acc += prev[__double2loint(en.y)] * en.x;
  }
  return acc;
}

// The reference's dot product is dense. Dropping its zero terms changes nothing but this: an infinity or a NaN
// in the previous row (ScaleMatrix multiplies a whole row by 2^256 while ANY entry is below 2^-256, so in the
// underflow regime the large entries overflow) times a zero coefficient is NaN, and a NaN term makes the sum NaN.
// So for a row that is not tame the sparse product is kept and the sum is poisoned exactly when some non-finite
// entry of `prev` falls outside the column's non-zero rows: `nf_total` non-finite entries in all (counted once
// per row by the warp), of which the loop below meets `nf_seen`. (finite * 0 = +-0 leaves every sum as it is.)
__device__ __forceinline__ int warp_count_nonfinite(const double* prev, int n, int lane) {
  int c = 0;
  for (int k = lane; k < n; k += 32) c += !(fabs(prev[k]) <= DBL_MAX);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  return c;
}
__device__ double csc_dot_dense(const CscDev& m, const double* prev, int j, int nf_total) {
  double acc = 0.0;
  int nf_seen = 0;
  const int e = m.colptr[j + 1];
  for (int p = m.colptr[j]; p < e; ++p) {
    const double2 en = __ldg(m.ent + p);
    const double x = prev[__double2loint(en.y)];
    nf_seen += !(fabs(x) <= DBL_MAX);
    acc += x * en.x;
  }
  if (nf_total > nf_seen) acc += __longlong_as_double(0x7ff8000000000000ll);  // inf * 0 somewhere in the dense sum
  return acc;
}

// ComputeJunctionForwardProbabilities (src/HMM.cpp:1107-1139). Most (state, site) pairs of a junction
// are silent - the state cannot emit at the site, its emission is 0 (FillJunctionEmission) - and their
// value is (dot product) * 0. While the previous row is tame (finite, <= 1e150; the matrices are checked
// on the host) that is exactly 0, so only the emitting states of a site get a dot product: 14 % of
// them at the benchmark shapes. A row that is not tame takes the full path.
__device__ int junction_forward(const FamilyDev& f, const CscDev& t_in, const CscDev& t_jj,
                                const int32_t* __restrict__ inds_rm, const int32_t* __restrict__ valid_off,
                                const int2* __restrict__ valid, int L, int Sj, const double* __restrict__ xe,
                                double*& prev, double*& cur, int count, bool& tame, double* fwd_out,
                                int32_t* cnt_out, int lane) {
  for (int i = 0; i < L; ++i) {
    const CscDev& m = (i == 0) ? t_in : t_jj;
    if (tame && f.skip_silent) {
      for (int j = lane; j < Sj; j += 32) cur[j] = 0.0;
      __syncwarp();
      const int e = valid_off[i + 1];
      for (int q = valid_off[i] + lane; q < e; q += 32) {
        const int2 sv = valid[q];
        cur[sv.x] = csc_dot(m, prev, sv.x) * xe[sv.y];
      }
    } else {
      const int nf = tame ? 0 : warp_count_nonfinite(prev, m.rows, lane);
      for (int j = lane; j < Sj; j += 32) {
        double acc = tame ? csc_dot(m, prev, j) : csc_dot_dense(m, prev, j, nf);
        const int idx = inds_rm[(size_t)i * Sj + j];
        acc *= (idx < 0) ? 0.0 : xe[idx];   // FillJunctionEmission: 0 where idx == -1
        cur[j] = acc;
      }
    }
    __syncwarp();
    count += warp_scale_row(cur, Sj, lane, tame);
    if (fwd_out)
      for (int j = lane; j < Sj; j += 32) fwd_out[(size_t)j * L + i] = cur[j];
    if (cnt_out && lane == 0) cnt_out[i] = count;
    double* tmp = prev; prev = cur; cur = tmp;
  }
  return count;
}

__global__ void __launch_bounds__(128) k_hmm(FamilyDev f, BatchDev b, HmmOut o, int tree0, int ntrees) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warps = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_warp = 2 * f.s_max + 5 * (f.g_max + (f.g_max + 1) / 2);
  double* base = reinterpret_cast<double*>(smem_raw) + (size_t)wid * per_warp;
  double* bufA = base;
  double* bufB = base + f.s_max;
  const int gstride = f.g_max + (f.g_max + 1) / 2;  // values + packed int counts
  double* e_vpad = base + 2 * f.s_max;
  double* e_vgerm = e_vpad + gstride;
  double* e_dgerm = e_vgerm + gstride;
  double* e_jgerm = e_dgerm + gstride;
  double* e_jpad = e_jgerm + gstride;

  for (int t = tree0 + blockIdx.x * warps + wid; t < tree0 + ntrees; t += gridDim.x * warps) {
    const double* xe = b.xe + (size_t)t * f.S;
    // emissions (src/PhyloHMM.cpp:94-114); counts reset per tree (:373-375)
    int c_v = fill_region(f.n_v, f.vpad_off, f.vpad_inds, xe, e_vpad, lane);
    c_v += fill_region(f.n_v, f.vgerm_off, f.vgerm_inds, xe, e_vgerm, lane);
    int c_d = 0;
    if (f.has_d) c_d = fill_region(f.n_d, f.dgerm_off, f.dgerm_inds, xe, e_dgerm, lane);
    int c_j = fill_region(f.n_j, f.jgerm_off, f.jgerm_inds, xe, e_jgerm, lane);
    c_j += fill_region(f.n_j, f.jpad_off, f.jpad_inds, xe, e_jpad, lane);

    // ComputeInitialForwardProbabilities (src/HMM.cpp:291-319)
    double* prev = bufA;
    double* cur = bufB;
    for (int g = lane; g < f.n_v; g += 32) {
      double v = f.v_gene_prob[g];
      v *= f.vpad_trans[g];
      v *= e_vpad[g];
      v *= f.v_trans_prod[g];
      v *= e_vgerm[g];
      prev[g] = v;
    }
    __syncwarp();
    bool tame;
    c_v += warp_scale_row(prev, f.n_v, lane, tame);
    if (o.vgerm_forward)
      for (int g = lane; g < f.n_v; g += 32) o.vgerm_forward[(size_t)t * f.n_v + g] = prev[g];

    // V-D (or V-J) junction
    int cnt = junction_forward(f, f.t_v_vd, f.t_vd, f.vd_inds_rm, f.vd_valid_off, f.vd_valid, f.vd_sites,
                               f.vd_states, xe, prev, cur, c_v, tame,
                               o.vd_forward ? o.vd_forward + (size_t)t * f.vd_sites * f.vd_states : nullptr,
                               o.vd_scalers ? o.vd_scalers + (size_t)t * f.vd_sites : nullptr, lane);
    if (f.has_d) {
      // ComputeGermlineForwardProbabilities for D (src/HMM.cpp:1160-1177; padding terms are ones)
      const int nf_d = tame ? 0 : warp_count_nonfinite(prev, f.t_vd_d.rows, lane);
      for (int g = lane; g < f.n_d; g += 32) {
        double v = tame ? csc_dot(f.t_vd_d, prev, g) : csc_dot_dense(f.t_vd_d, prev, g, nf_d);
        v *= e_dgerm[g];
        cur[g] = v;
      }
      __syncwarp();
      c_d += cnt + warp_scale_row(cur, f.n_d, lane, tame);
      if (o.dgerm_forward)
        for (int g = lane; g < f.n_d; g += 32) o.dgerm_forward[(size_t)t * f.n_d + g] = cur[g];
      { double* tmp = prev; prev = cur; cur = tmp; }
      cnt = junction_forward(f, f.t_d_dj, f.t_dj, f.dj_inds_rm, f.dj_valid_off, f.dj_valid, f.dj_sites,
                             f.dj_states, xe, prev, cur, c_d, tame,
                             o.dj_forward ? o.dj_forward + (size_t)t * f.dj_sites * f.dj_states : nullptr,
                             o.dj_scalers ? o.dj_scalers + (size_t)t * f.dj_sites : nullptr, lane);
    }
    const CscDev& t_out = f.has_d ? f.t_dj_j : f.t_vd_d;
    const int nf_j = tame ? 0 : warp_count_nonfinite(prev, t_out.rows, lane);
    for (int g = lane; g < f.n_j; g += 32) {
      double v = tame ? csc_dot(t_out, prev, g) : csc_dot_dense(t_out, prev, g, nf_j);
      v *= e_jgerm[g];
      v *= f.jpad_trans[g];
      v *= e_jpad[g];
      cur[g] = v;
    }
    __syncwarp();
    c_j += cnt + warp_scale_row(cur, f.n_j, lane);
    if (o.jgerm_forward)
      for (int g = lane; g < f.n_j; g += 32) o.jgerm_forward[(size_t)t * f.n_j + g] = cur[g];
    if (lane == 0) {
      double sum = 0.0;
      for (int g = 0; g < f.n_j; ++g) sum += cur[g];
      // HMM::LogLikelihood (src/HMM.cpp:352-353)
      b.loglik[t] = log(sum) - c_j * o.log_factor;
      if (o.germ_scalers) {
        o.germ_scalers[3 * t] = c_v;
        o.germ_scalers[3 * t + 1] = c_d;
        o.germ_scalers[3 * t + 2] = c_j;
      }
    }
    __syncwarp();
  }
}

size_t hmm_smem_bytes(const FamilyDev& f, int warps) {
  size_t per_warp = 2 * (size_t)f.s_max + 5 * ((size_t)f.g_max + (f.g_max + 1) / 2);
  return per_warp * warps * sizeof(double);
}

void launch_hmm(const FamilyDev& f, const BatchDev& b, const HmmOut& o, int tree0, int ntrees, cudaStream_t s) {
  const int warps = 4;
  size_t smem = hmm_smem_bytes(f, warps);
  ensure_dynamic_smem(reinterpret_cast<const void*>(&k_hmm), smem);
  int blocks = (ntrees + warps - 1) / warps;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_hmm<<<blocks, warps * 32, smem, s>>>(f, b, o, tree0, ntrees);
}

// ----------------------------------------------------------------- k_sample ---
// The random draws of HMM::SampleNaiveSequence (src/HMM.cpp:323-341, 1222-1353) for every tree of a scored batch,
// from the forward tables k_hmm left on the device: the J gene from jgerm_forward, then backwards through the D-J
// junction sites, the D gene, the V-D junction sites and the V gene, every draw a std::discrete_distribution over
// (transition column of the state drawn before) x (forward row). What the reference's standard library computes is
// reproduced operation by operation (libstdc++ bits/random.tcc): the weights summed left to right from 0.0, every
// weight divided by the sum, a running sum of those with the last entry forced to 1.0, u from
// generate_canonical<double, 53> = (x0 + x1 * 2^32) / 2^64 over two raw outputs of the generator (nextafter(1, 0) if
// that rounds to 1), and std::lower_bound = the first entry not below u. The raw std::mt19937 outputs come from the
// host, which advances the one generator of the run in file order (a fixed number of outputs per row, see
// rng_outputs_per_row in host/phylo_hmm.cpp); the strings of the annotation are put together there from the indices.
// The transition matrices are CSC (rows ascending) where the host multiplies dense columns: a zero entry adds 0.0 to
// the sum and leaves the running sum where it is, so the arithmetic is the same - as long as the forward row is
// finite. A row with a non-finite entry (0 x inf = NaN in the dense product) or a sum that is not a positive finite
// number is flagged and its batch is sampled by the host code instead.
// One warp per tree: the lanes form the weights of a draw in shared memory, lane 0 does the sequential sums.
struct SampleDev {
  const double *fv, *fvd, *fd, *fdj, *fj;   // forward tables of the batch (k_hmm's layout)
  const uint32_t* words;                    // [B][n_words] raw generator outputs, in draw order
  int32_t n_words;
  int32_t* idx;                             // [B][n_idx]: J gene | D-J states, site 0 first | D gene | V-D states | V gene
  int32_t n_idx;
  uint8_t* flag;                            // [B] 1: leave the tree to the host sampler
};

__device__ __forceinline__ double canonical_u(uint32_t x0, uint32_t x1) {
  const double sum = __dadd_rn((double)x0, __dmul_rn((double)x1, 4294967296.0));
  const double u = __dmul_rn(sum, 0x1p-64);
  return u >= 1.0 ? 0x1.fffffffffffffp-1 : u;
}

// one draw over n states with weights w[j] = (m ? m(j, col) : 1) * fwd[j * fstride]; returns the index to every lane
__device__ int sample_draw(const CscDev* m, int col, const double* __restrict__ fwd, int fstride, int n,
                           double* sw, int* sr, const uint32_t*& words, bool& bad, int lane) {
  if (n < 2) return 0;   // discrete_distribution keeps no weights then and draws nothing
  bool nonfinite = false;
  for (int j = lane; j < n; j += 32) nonfinite |= !(fabs(fwd[(size_t)j * fstride]) <= DBL_MAX);
  int cnt;
  if (m) {
    const int p0 = m->colptr[col];
    cnt = m->colptr[col + 1] - p0;
    for (int p = lane; p < cnt; p += 32) {
      const double2 en = __ldg(m->ent + p0 + p);
      const int row = __double2loint(en.y);
      sw[p] = __dmul_rn(en.x, fwd[(size_t)row * fstride]);
      sr[p] = row;
    }
  } else {
    cnt = n;
    for (int p = lane; p < n; p += 32) {
      sw[p] = fwd[(size_t)p * fstride];
      sr[p] = p;
    }
  }
  if (__any_sync(0xffffffffu, nonfinite)) bad = true;
  __syncwarp();
  int k = n - 1;
  // the sum is a chain (left to right, as std::accumulate adds), the quotients are not: the lanes divide in parallel
  // and lane 0's second chain is additions only
  double sum = 0.0;
  if (lane == 0) {
    for (int p = 0; p < cnt; ++p) sum = __dadd_rn(sum, sw[p]);
    if (!(sum > 0.0) || !(sum <= DBL_MAX)) bad = true;
  }
  sum = __shfl_sync(0xffffffffu, sum, 0);
  const bool sum_ok = __shfl_sync(0xffffffffu, (int)bad, 0) == 0;
  if (sum_ok)
    for (int p = lane; p < cnt; p += 32) sw[p] = __ddiv_rn(sw[p], sum);
  __syncwarp();
  if (lane == 0 && sum_ok) {
    const double u = canonical_u(words[0], words[1]);
    if (u == 0.0) {
      k = 0;
    } else {
      double cp = 0.0;
      for (int p = 0; p < cnt; ++p) {
        cp = __dadd_rn(cp, sw[p]);
        const double c = sr[p] == n - 1 ? 1.0 : cp;
        if (!(c < u)) {
          k = sr[p];
          break;
        }
      }
    }
  }
  words += 2;
  bad = __shfl_sync(0xffffffffu, (int)bad, 0) != 0 || bad;
  k = __shfl_sync(0xffffffffu, k, 0);
  __syncwarp();
  return k;
}

__global__ void __launch_bounds__(128) k_sample(FamilyDev f, SampleDev sd, int ntrees) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warps = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cap = max(max(f.s_max, f.g_max), 2);
  double* sw = reinterpret_cast<double*>(smem_raw) + (size_t)wid * cap;
  int* sr = reinterpret_cast<int*>(reinterpret_cast<double*>(smem_raw) + (size_t)warps * cap) + (size_t)wid * cap;
  for (int t = blockIdx.x * warps + wid; t < ntrees; t += gridDim.x * warps) {
    const uint32_t* words = sd.words + (size_t)t * sd.n_words;
    int32_t* idx = sd.idx + (size_t)t * sd.n_idx;
    bool bad = false;
    int at = 0;
    // the J gene (SampleInitialState)
    int germ = sample_draw(nullptr, 0, sd.fj + (size_t)t * f.n_j, 1, f.n_j, sw, sr, words, bad, lane);
    if (lane == 0) idx[at] = germ;
    ++at;
    // junctions backwards from the gene to their right, then the gene to their left (SampleJunctionStates,
    // SampleGermlineState); light chains have one junction and no D gene
    for (int pass = f.has_d ? 0 : 1; pass < 2; ++pass) {
      const bool dj = pass == 0;
      const int L = dj ? f.dj_sites : f.vd_sites, Sj = dj ? f.dj_states : f.vd_states;
      const double* fw = dj ? sd.fdj + (size_t)t * L * Sj : sd.fvd + (size_t)t * L * Sj;
      const CscDev& t_out = dj ? f.t_dj_j : f.t_vd_d;   // junction -> the gene to its right
      const CscDev& t_jj = dj ? f.t_dj : f.t_vd;
      int prev = germ;
      for (int i = L - 1; i >= 0; --i) {
        prev = sample_draw(i == L - 1 ? &t_out : &t_jj, prev, fw + i, L, Sj, sw, sr, words, bad, lane);
        if (lane == 0) idx[at + i] = prev;
      }
      at += L;
      // the gene to the left of the junction, given the junction's first state
      const int n_g = dj ? f.n_d : f.n_v;
      const double* fg = dj ? sd.fd + (size_t)t * n_g : sd.fv + (size_t)t * n_g;
      germ = sample_draw(dj ? &f.t_d_dj : &f.t_v_vd, prev, fg, 1, n_g, sw, sr, words, bad, lane);
      if (lane == 0) idx[at] = germ;
      ++at;
    }
    if (lane == 0) sd.flag[t] = bad ? 1 : 0;
  }
}

void launch_sample(const FamilyDev& f, const double* fv, const double* fvd, const double* fd, const double* fdj,
                   const double* fj, const uint32_t* words, int n_words, int32_t* idx, int n_idx, uint8_t* flag,
                   int ntrees, cudaStream_t s) {
  const int warps = 4;
  const int cap = std::max(std::max(f.s_max, f.g_max), 2);
  const size_t smem = (size_t)warps * cap * (sizeof(double) + sizeof(int));
  ensure_dynamic_smem(reinterpret_cast<const void*>(&k_sample), smem);
  SampleDev sd{fv, fvd, fd, fdj, fj, words, n_words, idx, n_idx, flag};
  int blocks = std::min((ntrees + warps - 1) / warps, 148 * 8);
  k_sample<<<std::max(blocks, 1), warps * 32, smem, s>>>(f, sd, ntrees);
}

// ------------------------------------------------------------- diagnostics ---
// Register-resident DFMA loop: the FP64 denominator for the on-chip pruning regime
// (MEASURED_PEAKS.json has no FP64 entry). 8 independent chains per thread.
__global__ void __launch_bounds__(256) k_fp64_peak(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
      x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
  }
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

double fp64_peak_flops_per_launch(double* scratch, int blocks, int iters, cudaStream_t s) {
  k_fp64_peak<<<blocks, 256, 0, s>>>(scratch, iters, 0.999999, 1e-9);
  return 2.0 * 64.0 * iters * 256.0 * blocks;
}

}  // namespace lhb
