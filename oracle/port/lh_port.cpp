// TEST INFRASTRUCTURE ONLY (oracle/). Not part of the product path: only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load this. The product (linearham_b200/) never links or calls it.
//
// Scalar CPU restatement of the reference's phylo-HMM likelihood path, taking
// the same flat inputs as the device C ABI (include/linearham_b200.h):
//   Gamma categories   libpll gamma.c:27-292
//   GTR eigen system   libpll models.c:24-331
//   P-matrices         libpll core_pmatrix.c:160-237
//   partials + scalers libpll core_partials.c:24-46, 575-663
//   edge log-lik       libpll core_likelihood_avx.c:1130-1260
//   naive correction   linearham src/PhyloHMM.cpp:228-237
//   emission fill      linearham src/PhyloHMM.cpp:158-215
//   scaled forward     linearham src/HMM.cpp:254-319, 1107-1177, 345-354
//   ScaleMatrix        linearham src/utils.cpp:135-144
// Parity of this restatement is pinned against the reference compiled in
// place (oracle/_ref, built by oracle/refbuild/Makefile) in
// tests/test_oracle_port.py and through the golden vectors in tests/golden/.
#include <cmath>
#include <cstdlib>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/linearham_b200.h"

namespace {

const double kScaleFactor = std::pow(2.0, 256);       // utils.hpp:22, pll.h:89-93
const double kScaleThreshold = 1.0 / kScaleFactor;

// ---------------------------------------------------------------- gamma.c ---

// gamma.c:27-95 (AS32 incomplete gamma ratio)
double incomplete_gamma(double x, double alpha, double ln_gamma_alpha) {
  const double accurate = 1e-8, overflow = 1e30;
  double p = alpha, g = ln_gamma_alpha;
  if (x == 0) return 0;
  if (x < 0 || p <= 0) return -1;
  double factor = std::exp(p * std::log(x) - x - g);
  if (!(x > 1 && x >= p)) {
    // series expansion
    double gin = 1, term = 1, rn = p;
    do {
      rn++;
      term *= x / rn;
      gin += term;
    } while (term > accurate);
    gin *= factor / p;  // note: gin * (factor / p), as gamma.c:61
    return gin;
  }
  // continued fraction
  double pn[6];
  double a = 1 - p, b = a + x + 1, term = 0;
  pn[0] = 1; pn[1] = x; pn[2] = x + 1; pn[3] = x * b;
  double gin = pn[2] / pn[3];
  for (;;) {
    a++;
    b += 2;
    term++;
    double an = a * term;
    for (int i = 0; i < 2; i++) pn[i + 4] = b * pn[i + 2] - an * pn[i];
    if (pn[5] != 0) {
      double rn = pn[4] / pn[5];
      double dif = std::fabs(gin - rn);
      if (dif <= accurate && dif <= accurate * rn) break;
      gin = rn;
    }
    for (int i = 0; i < 4; i++) pn[i] = pn[i + 2];
    if (std::fabs(pn[4]) >= overflow)
      for (int i = 0; i < 4; i++) pn[i] /= overflow;
  }
  return 1 - factor * gin;
}

// gamma.c:96-131 (Algorithm 291)
double ln_gamma(double alpha) {
  double x = alpha, f = 0.0, z;
  if (x < 7.0) {
    f = 1.0;
    z = alpha - 1.0;
    while ((z = z + 1.0) < 7.0) f *= z;
    x = z;
    f = -std::log(f);
  }
  z = 1 / (x * x);
  return f + (x - 0.5) * std::log(x) - x + .918938533204673 +
         (((-.000595238095238 * z + .000793650793651) * z - .002777777777778) * z +
          .083333333333333) / x;
}

// gamma.c:133-158 (AS70)
double point_normal(double prob) {
  const double a0 = -.322232431088, a1 = -1, a2 = -.342242088547, a3 = -.0204231210245;
  const double a4 = -.453642210148e-4, b0 = .0993484626060, b1 = .588581570495;
  const double b2 = .531103462366, b3 = .103537752850, b4 = .0038560700634;
  double p = prob, p1 = (p < 0.5 ? p : 1 - p);
  if (p1 < 1e-20) return -9999;
  double y = std::sqrt(std::log(1 / (p1 * p1)));
  double z = y + ((((y * a4 + a3) * y + a2) * y + a1) * y + a0) /
                     ((((y * b4 + b3) * y + b2) * y + b1) * y + b0);
  return (p < 0.5 ? -z : z);
}

// gamma.c:160-218 (AS91)
double point_chi2(double prob, double v) {
  const double e = .5e-6, aa = .6931471805;
  double p = prob;
  if (p < .000002 || p > .999998 || v <= 0) return -1;
  double g = ln_gamma(v / 2);
  double xx = v / 2, c = xx - 1, ch;
  if (v < -1.24 * std::log(p)) {
    ch = std::pow((p * xx * std::exp(g + xx * aa)), 1 / xx);
    if (ch - e < 0) return ch;
  } else if (v > .32) {
    double x = point_normal(p);
    double p1 = 0.222222 / v;
    ch = v * std::pow((x * std::sqrt(p1) + 1 - p1), 3.0);
    if (ch > 2.2 * v + 6) ch = -2 * (std::log(1 - p) - c * std::log(.5 * ch) + g);
  } else {
    ch = 0.4;
    double a = std::log(1 - p), q;
    do {
      q = ch;
      double p1 = 1 + ch * (4.67 + ch);
      double p2 = ch * (6.73 + ch * (6.66 + ch));
      double t = -0.5 + (4.67 + 2 * ch) / p1 - (6.73 + ch * (13.32 + 3 * ch)) / p2;
      ch -= (1 - std::exp(a + g + .5 * ch + c * aa) * p2 / p1) / t;
    } while (std::fabs(q / ch - 1) - .01 > 0);
  }
  double q;
  do {
    q = ch;
    double p1 = .5 * ch;
    double t = incomplete_gamma(p1, xx, g);
    if (t < 0.0) return -1;
    double p2 = p - t;
    t = p2 * std::exp(xx * aa + g + p1 - c * std::log(ch));
    double b = t / ch, a = 0.5 * t - b * c;
    double s1 = (210 + a * (140 + a * (105 + a * (84 + a * (70 + 60 * a))))) / 420;
    double s2 = (420 + a * (735 + a * (966 + a * (1141 + 1278 * a)))) / 2520;
    double s3 = (210 + a * (462 + a * (707 + 932 * a))) / 2520;
    double s4 = (252 + a * (672 + 1182 * a) + c * (294 + a * (889 + 1740 * a))) / 5040;
    double s5 = (84 + 264 * a + c * (175 + 606 * a)) / 2520;
    double s6 = (120 + c * (346 + 127 * c)) / 5040;
    ch += t * (1 + 0.5 * t * s1 - b * c * (s1 - b * (s2 - b * (s3 - b * (s4 - b * (s5 - b * s6))))));
  } while (std::fabs(q / ch - 1) > e);
  return ch;
}

}  // namespace

// gamma.c:220-292, PLL_GAMMA_RATES_MEAN branch only (src/PhyloHMM.cpp:360,425)
extern "C" int port_gamma_cats(double alpha, int categories, double* out) {
  if (alpha < 0.02 || categories < 1) return 1;
  if (categories == 1) { out[0] = 1.0; return 0; }
  double factor = alpha / alpha * categories, beta = alpha;
  std::vector<double> gp(categories);
  double lnga1 = ln_gamma(alpha + 1);
  for (int i = 0; i < categories - 1; i++)
    gp[i] = point_chi2((i + 1.0) / categories, 2.0 * alpha) / (2.0 * beta);
  for (int i = 0; i < categories - 1; i++)
    gp[i] = incomplete_gamma(gp[i] * beta, alpha + 1, lnga1);
  out[0] = gp[0] * factor;
  out[categories - 1] = (1 - gp[categories - 2]) * factor;
  for (int i = 1; i < categories - 1; i++) out[i] = (gp[i] - gp[i - 1]) * factor;
  return 0;
}

namespace {

// ---------------------------------------------------------------- models.c ---

// Householder tridiagonalisation of a symmetric n x n matrix held in a[][]
// (models.c:100-178, the classic tred2), 0-based restatement.
void tridiagonalise(double a[4][4], int n, double* d, double* e) {
  for (int i = n - 1; i >= 1; i--) {
    int l = i;  // number of leading entries of column i above the diagonal
    double h = 0.0, scale = 0.0;
    if (l > 1) {
      for (int k = 0; k < l; k++) scale += std::fabs(a[k][i]);
      if (scale == 0.0) {
        e[i] = a[l - 1][i];
      } else {
        for (int k = 0; k < l; k++) {
          a[k][i] /= scale;
          h += a[k][i] * a[k][i];
        }
        double f = a[l - 1][i];
        double g = (f > 0) ? -std::sqrt(h) : std::sqrt(h);
        e[i] = scale * g;
        h -= f * g;
        a[l - 1][i] = f - g;
        f = 0.0;
        for (int j = 0; j < l; j++) {
          a[i][j] = a[j][i] / h;
          g = 0.0;
          for (int k = 0; k <= j; k++) g += a[k][j] * a[k][i];
          for (int k = j + 1; k < l; k++) g += a[j][k] * a[k][i];
          e[j] = g / h;
          f += e[j] * a[j][i];
        }
        double hh = f / (h + h);
        for (int j = 0; j < l; j++) {
          f = a[j][i];
          g = e[j] - hh * f;
          e[j] = g;
          for (int k = 0; k <= j; k++) a[k][j] -= (f * e[k] + g * a[k][i]);
        }
      }
    } else {
      e[i] = a[l - 1][i];
    }
    d[i] = h;
  }
  d[0] = 0.0;
  e[0] = 0.0;
  for (int i = 0; i < n; i++) {
    int l = i;
    if (d[i] != 0.0) {
      for (int j = 0; j < l; j++) {
        double g = 0.0;
        for (int k = 0; k < l; k++) g += a[k][i] * a[j][k];
        for (int k = 0; k < l; k++) a[j][k] -= g * a[i][k];
      }
    }
    d[i] = a[i][i];
    a[i][i] = 1.0;
    for (int j = 0; j < l; j++) a[i][j] = a[j][i] = 0.0;
  }
}

// Implicit-shift QL on the tridiagonal (d, e), accumulating into z
// (models.c:24-97, the classic tqli), 0-based restatement.
void ql_implicit(double* d, double* e, int n, double z[4][4]) {
  for (int i = 1; i < n; i++) e[i - 1] = e[i];
  e[n - 1] = 0.0;
  for (int l = 0; l < n; l++) {
    int m;
    do {
      for (m = l; m < n - 1; m++) {
        double dd = std::fabs(d[m]) + std::fabs(d[m + 1]);
        if (std::fabs(e[m]) + dd == dd) break;
      }
      if (m != l) {
        double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
        double r = std::sqrt((g * g) + 1.0);
        g = d[m] - d[l] + e[l] / (g + ((g < 0) ? -std::fabs(r) : std::fabs(r)));
        double s = 1.0, c = 1.0, p = 0.0;
        for (int i = m - 1; i >= l; i--) {
          double f = s * e[i];
          double b = c * e[i];
          if (std::fabs(f) >= std::fabs(g)) {
            c = g / f;
            r = std::sqrt((c * c) + 1.0);
            e[i + 1] = f * r;
            c *= (s = 1.0 / r);
          } else {
            s = f / g;
            r = std::sqrt((s * s) + 1.0);
            e[i + 1] = g * r;
            s *= (c = 1.0 / r);
          }
          g = d[i + 1] - p;
          r = (d[i] - g) * s + 2.0 * c * b;
          p = s * r;
          d[i + 1] = g + p;
          g = c * r - b;
          for (int k = 0; k < n; k++) {
            f = z[i + 1][k];
            z[i + 1][k] = s * z[i][k] + c * f;
            z[i][k] = c * z[i][k] - s * f;
          }
        }
        d[l] = d[l] - p;
        e[l] = g;
        e[m] = 0.0;
      }
    } while (m != l);
  }
}

struct GtrEigen {
  double evals[4], evecs[16], inv_evecs[16];
};

// create_ratematrix + pll_update_eigen (models.c:182-331)
void gtr_eigen(const double* er, const double* pi, GtrEigen* out) {
  double r[6];
  for (int i = 0; i < 6; i++) r[i] = er[i];
  if (r[5] > 0.0)
    for (int i = 0; i < 6; i++) r[i] /= r[5];
  double q[4][4];
  for (int i = 0; i < 4; i++) q[i][i] = 0;
  int k = 0;
  for (int i = 0; i < 4; i++)
    for (int j = i + 1; j < 4; j++) {
      double factor = r[k++];
      q[i][j] = q[j][i] = factor * std::sqrt(pi[i] * pi[j]);
      q[i][i] -= factor * pi[j];
      q[j][j] -= factor * pi[i];
    }
  double mean = 0;
  for (int i = 0; i < 4; i++) mean += pi[i] * (-q[i][i]);
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) q[i][j] /= mean;
  double d[4], e[4];
  tridiagonalise(q, 4, d, e);
  ql_implicit(d, e, 4, q);
  for (int i = 0; i < 4; i++) {
    out->evals[i] = d[i];
    for (int j = 0; j < 4; j++) out->evecs[i * 4 + j] = q[i][j];
  }
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) out->inv_evecs[i * 4 + j] = out->evecs[j * 4 + i];
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) out->inv_evecs[i * 4 + j] /= std::sqrt(pi[i]);
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) out->evecs[i * 4 + j] *= std::sqrt(pi[j]);
}

// core_pmatrix.c:160-237 (prop_invar == 0)
void pmatrix(const GtrEigen& g, double rate, double t, double* pmat) {
  if (!t) {
    for (int j = 0; j < 4; j++)
      for (int k = 0; k < 4; k++) pmat[j * 4 + k] = (j == k) ? 1 : 0;
    return;
  }
  double expd[4], temp[16];
  for (int j = 0; j < 4; j++) expd[j] = std::expm1(g.evals[j] * rate * t);
  for (int j = 0; j < 4; j++)
    for (int k = 0; k < 4; k++) temp[j * 4 + k] = g.inv_evecs[j * 4 + k] * expd[k];
  for (int j = 0; j < 4; j++)
    for (int k = 0; k < 4; k++) {
      double v = (j == k) ? 1.0 : 0;
      for (int m = 0; m < 4; m++) v += temp[j * 4 + m] * g.evecs[m * 4 + k];
      pmat[j * 4 + k] = v;
    }
}

// pll_set_tip_states / set_tipclv (pll.c:905-964) with pll_map_nt (maps.c:46-64)
// restricted to the xMSA alphabet ACGTN (codes 0..4).
inline void tip_clv(int code, int R, double* clv) {
  unsigned mask = (code == 4) ? 15u : (1u << code);
  for (int k = 0; k < R; k++)
    for (int j = 0; j < 4; j++) clv[k * 4 + j] = (mask >> j) & 1u ? 1.0 : 0.0;
}

// utils.cpp:135-144
int scale_row(double* m, int n) {
  int count = 0;
  for (;;) {
    bool any = false;
    for (int i = 0; i < n; i++)
      if (0 < m[i] && m[i] < kScaleThreshold) any = true;
    if (!any) break;
    for (int i = 0; i < n; i++) m[i] *= kScaleFactor;
    count++;
  }
  return count;
}

// src/PhyloHMM.cpp:158-193
void fill_germline_padding(int n, const int32_t* off, const int32_t* inds,
                           const double* xe, std::vector<double>& emission,
                           int& scaler_count) {
  emission.assign(n, 1.0);
  std::vector<int> counts(n, 0);
  int max_count = 0;
  for (int i = 0; i < n; i++) {
    for (int j = off[i]; j < off[i + 1]; j++) {
      emission[i] *= xe[inds[j]];
      counts[i] += scale_row(&emission[i], 1);
    }
    if (counts[i] > max_count) max_count = counts[i];
  }
  scaler_count += max_count;
  for (int i = 0; i < n; i++)
    emission[i] *= std::pow(kScaleFactor, max_count - counts[i]);
}

}  // namespace

// Column-major dense helpers matching Eigen's layout.
static inline double cm(const double* m, int rows, int i, int j) { return m[(size_t)j * rows + i]; }

// Score one sample; optional read-outs may be NULL.
static double port_score_one(const lh_family_desc_t& F, int R, const int32_t* children,
                             const double* blen, int root_child, double root_blen,
                             const double* er, const double* pi, const double* rates,
                             double* xe_out, double* vgerm_fwd_out, double* vd_fwd_out,
                             double* dgerm_fwd_out, double* dj_fwd_out,
                             double* jgerm_fwd_out, int32_t* germ_scalers_out,
                             int32_t* vd_scalers_out, int32_t* dj_scalers_out, double* srl_out = nullptr) {
  const int T = F.n_tips, S = F.n_sites, span = 4 * R, n_ops = T - 2;

  // ---- pruning: libptpll TraversalUpdate(FULL) + LogLikelihood ------------
  GtrEigen eig;
  gtr_eigen(er, pi, &eig);

  // LH_PORT_RATE_SCALERS=1 (experiments only, tests/ never set it): scale every rate category on its own count, the
  // way the CUDA kernels did before they took libpll's all-categories rule; used to find inputs that tell the two apart
  static const bool rate_scalers = getenv("LH_PORT_RATE_SCALERS") != nullptr;
  std::vector<double> clv((size_t)(2 * T - 2) * S * span);
  std::vector<unsigned> scaler((size_t)n_ops * S, 0u);
  std::vector<unsigned> rscaler(rate_scalers ? (size_t)n_ops * S * R : 0, 0u);
  for (int t = 0; t < T; t++)
    for (int n = 0; n < S; n++)
      tip_clv(F.tip_codes[(size_t)t * S + n], R, &clv[((size_t)t * S + n) * span]);

  std::vector<double> pl(16 * R), pr(16 * R);
  for (int op = 0; op < n_ops; op++) {
    int c1 = children[2 * op], c2 = children[2 * op + 1];
    for (int k = 0; k < R; k++) {
      pmatrix(eig, rates[k], blen[2 * op], &pl[16 * k]);
      pmatrix(eig, rates[k], blen[2 * op + 1], &pr[16 * k]);
    }
    const unsigned* ls = (c1 >= T) ? &scaler[(size_t)(c1 - T) * S] : nullptr;
    const unsigned* rs = (c2 >= T) ? &scaler[(size_t)(c2 - T) * S] : nullptr;
    unsigned* ps = &scaler[(size_t)op * S];
    // fill_parent_scaler, core_partials.c:24-46
    for (int n = 0; n < S; n++) ps[n] = (ls ? ls[n] : 0u) + (rs ? rs[n] : 0u);
    double* parent = &clv[(size_t)(T + op) * S * span];
    const double* left = &clv[(size_t)c1 * S * span];
    const double* right = &clv[(size_t)c2 * S * span];
    // core_partials.c:606-661 (per-site scaling mode)
    for (int n = 0; n < S; n++) {
      bool site_scale = true;
      for (int k = 0; k < R; k++) {
        bool rate_scale = true;
        for (int i = 0; i < 4; i++) {
          double ta = 0, tb = 0;
          for (int j = 0; j < 4; j++) {
            ta += pl[16 * k + 4 * i + j] * left[(size_t)n * span + 4 * k + j];
            tb += pr[16 * k + 4 * i + j] * right[(size_t)n * span + 4 * k + j];
          }
          double v = ta * tb;
          parent[(size_t)n * span + 4 * k + i] = v;
          rate_scale = rate_scale && (v < kScaleThreshold);
        }
        site_scale = site_scale && rate_scale;
      }
      if (site_scale && !rate_scalers) {
        for (int i = 0; i < span; i++) parent[(size_t)n * span + i] *= kScaleFactor;
        ps[n] += 1;
      }
      if (rate_scalers)
        for (int k = 0; k < R; k++) {
          unsigned& c = rscaler[((size_t)op * S + n) * R + k];
          c = (c1 >= T ? rscaler[((size_t)(c1 - T) * S + n) * R + k] : 0u) + (c2 >= T ? rscaler[((size_t)(c2 - T) * S + n) * R + k] : 0u);
          bool small = true;
          for (int i = 0; i < 4; i++) small = small && parent[(size_t)n * span + 4 * k + i] < kScaleThreshold;
          if (small) {
            for (int i = 0; i < 4; i++) parent[(size_t)n * span + 4 * k + i] *= kScaleFactor;
            c += 1;
          }
        }
    }
  }

  // edge log-likelihood at (virtual root, root_child), core_likelihood_avx.c:1130-1260
  std::vector<double> xe(S);
  {
    std::vector<double> pm(16 * R);
    for (int k = 0; k < R; k++) pmatrix(eig, rates[k], root_blen, &pm[16 * k]);
    const int root = T + n_ops - 1;
    const double* clvp = &clv[(size_t)root * S * span];
    const double* clvc = &clv[(size_t)root_child * S * span];
    const unsigned* sp = &scaler[(size_t)(n_ops - 1) * S];
    const unsigned* sc = (root_child >= T) ? &scaler[(size_t)(root_child - T) * S] : nullptr;
    const double weight = 1.0 / R;  // pll.c:746-747
    std::vector<double> rate_term(R);
    for (int n = 0; n < S; n++) {
      double terma = 0;
      for (int k = 0; k < R; k++) {
        double terma_r = 0;
        for (int i = 0; i < 4; i++) {
          double termb = 0;
          for (int j = 0; j < 4; j++)
            termb += pm[16 * k + 4 * i + j] * clvc[(size_t)n * span + 4 * k + j];
          terma_r += pi[i] * termb * clvp[(size_t)n * span + 4 * k + i];
        }
        if (terma_r > 0. && !rate_scalers) terma += terma_r * weight;
        rate_term[k] = terma_r;
      }
      unsigned cnt = sp[n] + (sc ? sc[n] : 0u);
      if (rate_scalers) {  // sum the categories at the smallest count (exact: powers of two)
        double acc = 0.0;
        unsigned acc_cnt = 0;
        for (int k = 0; k < R; k++) {
          double term = rate_term[k];
          unsigned c = rscaler[((size_t)(n_ops - 1) * S + n) * R + k] + (root_child >= T ? rscaler[((size_t)(root_child - T) * S + n) * R + k] : 0u);
          if (!(term > 0.)) continue;
          term *= weight;
          if (acc == 0.0) acc_cnt = c;
          while (c > acc_cnt) { term *= kScaleThreshold; --c; if (term == 0.0) c = acc_cnt; }
          while (acc_cnt > c) { acc *= kScaleThreshold; --acc_cnt; if (acc == 0.0) acc_cnt = c; }
          acc += term;
        }
        terma = acc;
        cnt = acc_cnt;
      }
      double site_lk = std::log(terma);
      // per-category read-out: what scripts/run_bootstrap_asr_ess.R:84-87 recomputes with phylomd
      // (phylo.likelihood on the rate-scaled tree / pi[naive base]); the site's scaler count is shared
      // by its categories (core_partials.c:648-659)
      if (srl_out)
        for (int k = 0; k < R; k++) {
          double v = std::log(rate_term[k]);
          if (cnt) v += cnt * std::log(kScaleThreshold);
          if (F.tip_codes[n] != 4) v -= std::log(pi[F.tip_codes[n]]);
          srl_out[(size_t)k * S + n] = rate_term[k] > 0. ? v : -INFINITY;
        }
      if (cnt) site_lk += cnt * std::log(kScaleThreshold);
      // naive correction, src/PhyloHMM.cpp:228-237 (row 0 of the xMSA is the naive)
      int nb = F.tip_codes[n];
      if (nb != 4) site_lk -= std::log(pi[nb]);
      // xmsa_emission_.array().exp() (src/PhyloHMM.cpp:239) is Eigen's packet exp (pexp_double, SSE2 Packet2d in the
      // reference's build): 2^n * exp(g) with n = floor(x log2(e) + 1/2), which is 0 once n <= -1023; the odd last
      // element of the vector goes through libm
      const bool packet = n < (S & ~1);
      xe[n] = (packet && std::floor(site_lk * 1.4426950408889634073599 + 0.5) <= -1023.0) ? 0.0 : std::exp(site_lk);
    }
  }
  if (xe_out) std::memcpy(xe_out, xe.data(), sizeof(double) * S);

  // ---- emissions, src/PhyloHMM.cpp:94-114 ---------------------------------
  int c_v = 0, c_d = 0, c_j = 0;
  std::vector<double> vpadE, vgermE, dgermE, jgermE, jpadE;
  fill_germline_padding(F.n_v, F.vpad_off, F.vpad_inds, xe.data(), vpadE, c_v);
  fill_germline_padding(F.n_v, F.vgerm_off, F.vgerm_inds, xe.data(), vgermE, c_v);
  if (F.has_d) fill_germline_padding(F.n_d, F.dgerm_off, F.dgerm_inds, xe.data(), dgermE, c_d);
  fill_germline_padding(F.n_j, F.jgerm_off, F.jgerm_inds, xe.data(), jgermE, c_j);
  fill_germline_padding(F.n_j, F.jpad_off, F.jpad_inds, xe.data(), jpadE, c_j);

  auto junction_emission = [&](const int32_t* inds, int L, int Sj, int i, int j) {
    int idx = inds[(size_t)j * L + i];
    return idx == -1 ? 0.0 : xe[idx];
  };

  // ---- forward, src/HMM.cpp:254-319 ---------------------------------------
  std::vector<double> fv(F.n_v);
  for (int g = 0; g < F.n_v; g++) {
    double f = F.v_gene_prob[g];
    f *= F.vpad_trans[g];
    f *= vpadE[g];
    f *= F.v_trans_prod[g];
    f *= vgermE[g];
    fv[g] = f;
  }
  c_v += scale_row(fv.data(), F.n_v);
  if (vgerm_fwd_out) std::memcpy(vgerm_fwd_out, fv.data(), sizeof(double) * F.n_v);

  // src/HMM.cpp:1107-1139
  auto junction = [&](const std::vector<double>& germ_fwd, int germ_count, int n_from,
                      const double* t_in, const double* t_jj, const int32_t* inds, int L,
                      int Sj, std::vector<double>& fwd, std::vector<int>& counts) {
    fwd.assign((size_t)L * Sj, 0.0);  // row-major here: fwd[i*Sj + j]
    counts.assign(L, 0);
    for (int i = 0; i < L; i++) {
      double* row = &fwd[(size_t)i * Sj];
      int prev;
      if (i == 0) {
        for (int j = 0; j < Sj; j++) {
          double acc = 0;
          for (int k = 0; k < n_from; k++) acc += germ_fwd[k] * cm(t_in, n_from, k, j);
          row[j] = acc;
        }
        prev = germ_count;
      } else {
        const double* prow = &fwd[(size_t)(i - 1) * Sj];
        for (int j = 0; j < Sj; j++) {
          double acc = 0;
          for (int k = 0; k < Sj; k++) acc += prow[k] * cm(t_jj, Sj, k, j);
          row[j] = acc;
        }
        prev = counts[i - 1];
      }
      for (int j = 0; j < Sj; j++) row[j] *= junction_emission(inds, L, Sj, i, j);
      counts[i] = prev + scale_row(row, Sj);
    }
  };
  // src/HMM.cpp:1160-1177
  auto germline = [&](const std::vector<double>& jfwd, const std::vector<int>& jcounts, int L,
                      int Sj, const double* t_out, int n_to, const std::vector<double>& germE,
                      const double* padT, const std::vector<double>* padE,
                      std::vector<double>& out, int& count) {
    out.assign(n_to, 0.0);
    const double* last = &jfwd[(size_t)(L - 1) * Sj];
    for (int g = 0; g < n_to; g++) {
      double acc = 0;
      for (int k = 0; k < Sj; k++) acc += last[k] * cm(t_out, Sj, k, g);
      acc *= germE[g];
      acc *= padT ? padT[g] : 1.0;
      acc *= padE ? (*padE)[g] : 1.0;
      out[g] = acc;
    }
    count += jcounts.back() + scale_row(out.data(), n_to);
  };

  std::vector<double> vd_fwd, dj_fwd, fd, fj;
  std::vector<int> vd_counts, dj_counts;
  junction(fv, c_v, F.n_v, F.t_v_vd, F.t_vd, F.vd_inds, F.vd_sites, F.vd_states, vd_fwd, vd_counts);
  if (F.has_d) {
    germline(vd_fwd, vd_counts, F.vd_sites, F.vd_states, F.t_vd_d, F.n_d, dgermE, nullptr, nullptr, fd, c_d);
    junction(fd, c_d, F.n_d, F.t_d_dj, F.t_dj, F.dj_inds, F.dj_sites, F.dj_states, dj_fwd, dj_counts);
    germline(dj_fwd, dj_counts, F.dj_sites, F.dj_states, F.t_dj_j, F.n_j, jgermE, F.jpad_trans, &jpadE, fj, c_j);
  } else {
    germline(vd_fwd, vd_counts, F.vd_sites, F.vd_states, F.t_vd_d, F.n_j, jgermE, F.jpad_trans, &jpadE, fj, c_j);
  }

  auto to_colmajor = [](const std::vector<double>& rm, int L, int Sj, double* out) {
    for (int i = 0; i < L; i++)
      for (int j = 0; j < Sj; j++) out[(size_t)j * L + i] = rm[(size_t)i * Sj + j];
  };
  if (vd_fwd_out) to_colmajor(vd_fwd, F.vd_sites, F.vd_states, vd_fwd_out);
  if (dj_fwd_out && F.has_d) to_colmajor(dj_fwd, F.dj_sites, F.dj_states, dj_fwd_out);
  if (dgerm_fwd_out && F.has_d) std::memcpy(dgerm_fwd_out, fd.data(), sizeof(double) * F.n_d);
  if (jgerm_fwd_out) std::memcpy(jgerm_fwd_out, fj.data(), sizeof(double) * F.n_j);
  if (germ_scalers_out) { germ_scalers_out[0] = c_v; germ_scalers_out[1] = c_d; germ_scalers_out[2] = c_j; }
  if (vd_scalers_out) for (int i = 0; i < F.vd_sites; i++) vd_scalers_out[i] = vd_counts[i];
  if (dj_scalers_out && F.has_d) for (int i = 0; i < F.dj_sites; i++) dj_scalers_out[i] = dj_counts[i];

  // src/HMM.cpp:345-354
  double sum = 0;
  for (int g = 0; g < F.n_j; g++) sum += fj[g];
  return std::log(sum) - c_j * std::log(kScaleFactor);
}

extern "C" int port_score_batch(const lh_family_desc_t* F, const lh_tree_batch_t* tb,
                                double* loglik, lh_outputs_t* o) {
  const int T = F->n_tips, n_ops = T - 2, R = tb->n_rates, S = F->n_sites;
  const size_t vd = (size_t)F->vd_sites * F->vd_states, dj = (size_t)F->dj_sites * F->dj_states;
  for (int b = 0; b < tb->n_trees; b++) {
    loglik[b] = port_score_one(
        *F, R, tb->children + (size_t)b * n_ops * 2, tb->blen + (size_t)b * n_ops * 2,
        tb->root_child[b], tb->root_blen[b], tb->er + 6 * b, tb->pi + 4 * b,
        tb->rates + (size_t)R * b,
        (o && o->xmsa_emission) ? o->xmsa_emission + (size_t)b * S : nullptr,
        (o && o->vgerm_forward) ? o->vgerm_forward + (size_t)b * F->n_v : nullptr,
        (o && o->vd_forward) ? o->vd_forward + b * vd : nullptr,
        (o && o->dgerm_forward) ? o->dgerm_forward + (size_t)b * F->n_d : nullptr,
        (o && o->dj_forward) ? o->dj_forward + b * dj : nullptr,
        (o && o->jgerm_forward) ? o->jgerm_forward + (size_t)b * F->n_j : nullptr,
        (o && o->germ_scalers) ? o->germ_scalers + 3 * b : nullptr,
        (o && o->vd_scalers) ? o->vd_scalers + (size_t)b * F->vd_sites : nullptr,
        (o && o->dj_scalers) ? o->dj_scalers + (size_t)b * F->dj_sites : nullptr,
        (o && o->site_rate_loglik) ? o->site_rate_loglik + (size_t)b * R * S : nullptr);
  }
  return 0;
}
