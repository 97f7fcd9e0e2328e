// Discrete-Gamma rate categories on the host (one call per tree sample; ~1 us).
//
// The reference takes these from libpll's pll_compute_gamma_cats(alpha, R, out,
// PLL_GAMMA_RATES_MEAN) (gamma.c:220-292, called at src/PhyloHMM.cpp:360,425). That
// routine stops its series at 1e-8 / 5e-7 (gamma.c:41,169), so its output carries a
// truncation error of ~1e-11..1e-8 that enters every P-matrix: a mathematically better
// quantile routine would move log-likelihoods by more than the 1e-9 parity bar. The
// functions below therefore evaluate the same published algorithms (Lau's Algorithm 291,
// Bhattacharjee's AS32, Best & Roberts' AS91, Odeh & Evans' AS70) with the same stopping
// rules and the same order of floating-point operations as gamma.c. Compiled with
// -ffp-contract=off so no FMA contraction changes the rounding.
#include <cmath>
#include <vector>

#include "../../../include/linearham_b200_host.h"
#include "host_error.hpp"

namespace lhb {
namespace {

// ln Gamma(a), a > 0: Stirling series after shifting a up to >= 7 (gamma.c:96-131)
double lgamma_stirling(double a) {
  double x = a, shift = 0.0;
  if (x < 7.0) {
    double prod = 1.0, z = a - 1.0;
    while ((z = z + 1.0) < 7.0) prod *= z;
    x = z;
    shift = -std::log(prod);
  }
  const double z = 1 / (x * x);
  return shift + (x - 0.5) * std::log(x) - x + .918938533204673 +
         (((-.000595238095238 * z + .000793650793651) * z - .002777777777778) * z + .083333333333333) / x;
}

// regularised lower incomplete gamma P(shape, x) (gamma.c:27-95); lg = ln Gamma(shape)
double gamma_p(double x, double shape, double lg) {
  const double tol = 1e-8, big = 1e30;
  if (x == 0) return 0;
  if (x < 0 || shape <= 0) return -1;
  const double front = std::exp(shape * std::log(x) - x - lg);
  if (!(x > 1 && x >= shape)) {
    double sum = 1, term = 1, n = shape;
    do {
      n++;
      term *= x / n;
      sum += term;
    } while (term > tol);
    sum *= front / shape;
    return sum;
  }
  // continued fraction, evaluated by the forward recurrence on two convergent pairs
  double w[6];
  double a = 1 - shape, b = a + x + 1, count = 0;
  w[0] = 1;
  w[1] = x;
  w[2] = x + 1;
  w[3] = x * b;
  double cf = w[2] / w[3];
  for (;;) {
    a++;
    b += 2;
    count++;
    const double an = a * count;
    w[4] = b * w[2] - an * w[0];
    w[5] = b * w[3] - an * w[1];
    if (w[5] != 0) {
      const double next = w[4] / w[5];
      const double gap = std::fabs(cf - next);
      if (gap <= tol && gap <= tol * next) break;
      cf = next;
    }
    w[0] = w[2];
    w[1] = w[3];
    w[2] = w[4];
    w[3] = w[5];
    if (std::fabs(w[4]) >= big) {
      w[0] /= big;
      w[1] /= big;
      w[2] /= big;
      w[3] /= big;
    }
  }
  return 1 - front * cf;
}

// standard normal quantile (gamma.c:133-158)
double normal_quantile(double p) {
  const double a0 = -.322232431088, a1 = -1, a2 = -.342242088547, a3 = -.0204231210245;
  const double a4 = -.453642210148e-4, b0 = .0993484626060, b1 = .588581570495;
  const double b2 = .531103462366, b3 = .103537752850, b4 = .0038560700634;
  const double tail = (p < 0.5 ? p : 1 - p);
  if (tail < 1e-20) return -9999;
  const double y = std::sqrt(std::log(1 / (tail * tail)));
  const double z = y + ((((y * a4 + a3) * y + a2) * y + a1) * y + a0) /
                           ((((y * b4 + b3) * y + b2) * y + b1) * y + b0);
  return (p < 0.5 ? -z : z);
}

// chi-square quantile with v degrees of freedom (gamma.c:160-218)
double chi2_quantile(double p, double v) {
  const double eps = .5e-6, ln2 = .6931471805;
  if (p < .000002 || p > .999998 || v <= 0) return -1;
  const double lg = lgamma_stirling(v / 2);
  const double half = v / 2, c = half - 1;
  double ch;
  if (v < -1.24 * std::log(p)) {
    ch = std::pow((p * half * std::exp(lg + half * ln2)), 1 / half);
    if (ch - eps < 0) return ch;
  } else if (v > .32) {
    const double x = normal_quantile(p);
    const double p1 = 0.222222 / v;
    ch = v * std::pow((x * std::sqrt(p1) + 1 - p1), 3.0);
    if (ch > 2.2 * v + 6) ch = -2 * (std::log(1 - p) - c * std::log(.5 * ch) + lg);
  } else {
    ch = 0.4;
    const double a = std::log(1 - p);
    double prev;
    do {
      prev = ch;
      const double p1 = 1 + ch * (4.67 + ch);
      const double p2 = ch * (6.73 + ch * (6.66 + ch));
      const double t = -0.5 + (4.67 + 2 * ch) / p1 - (6.73 + ch * (13.32 + 3 * ch)) / p2;
      ch -= (1 - std::exp(a + lg + .5 * ch + c * ln2) * p2 / p1) / t;
    } while (std::fabs(prev / ch - 1) - .01 > 0);
  }
  double prev;
  do {  // seven-term Taylor refinement
    prev = ch;
    const double p1 = .5 * ch;
    double t = gamma_p(p1, half, lg);
    if (t < 0.0) return -1;
    const double p2 = p - t;
    t = p2 * std::exp(half * ln2 + lg + p1 - c * std::log(ch));
    const double b = t / ch, a = 0.5 * t - b * c;
    const double s1 = (210 + a * (140 + a * (105 + a * (84 + a * (70 + 60 * a))))) / 420;
    const double s2 = (420 + a * (735 + a * (966 + a * (1141 + 1278 * a)))) / 2520;
    const double s3 = (210 + a * (462 + a * (707 + 932 * a))) / 2520;
    const double s4 = (252 + a * (672 + 1182 * a) + c * (294 + a * (889 + 1740 * a))) / 5040;
    const double s5 = (84 + 264 * a + c * (175 + 606 * a)) / 2520;
    const double s6 = (120 + c * (346 + 127 * c)) / 5040;
    ch += t * (1 + 0.5 * t * s1 - b * c * (s1 - b * (s2 - b * (s3 - b * (s4 - b * (s5 - b * s6))))));
  } while (std::fabs(prev / ch - 1) > eps);
  return ch;
}

}  // namespace

// Mean rate of each of n equal-probability slices of Gamma(alpha, alpha).
int gamma_mean_rates(double alpha, int n, double* out) {
  if (!(alpha >= 0.02)) return host_fail("alpha %g is below the 0.02 the Gamma routine accepts", alpha);
  if (n < 1) return host_fail("need at least one rate category");
  if (n == 1) {
    out[0] = 1.0;
    return 0;
  }
  const double scale = alpha / alpha * n, beta = alpha;
  std::vector<double> cut(n);
  const double lg1 = lgamma_stirling(alpha + 1);
  for (int i = 0; i < n - 1; i++) cut[i] = chi2_quantile((i + 1.0) / n, 2.0 * alpha) / (2.0 * beta);
  for (int i = 0; i < n - 1; i++) cut[i] = gamma_p(cut[i] * beta, alpha + 1, lg1);
  out[0] = cut[0] * scale;
  out[n - 1] = (1 - cut[n - 2]) * scale;
  for (int i = 1; i < n - 1; i++) out[i] = (cut[i] - cut[i - 1]) * scale;
  return 0;
}

}  // namespace lhb

extern "C" int lhh_gamma_cats(double alpha, int32_t n_rates, double* out) {
  return lhb::gamma_mean_rates(alpha, n_rates, out);
}

extern "C" int lhh_gamma_cats_batch(int32_t n, const double* alpha, int32_t n_rates, double* out) {
  for (int i = 0; i < n; ++i)
    if (lhb::gamma_mean_rates(alpha[i], n_rates, out + (size_t)i * n_rates)) return 1;
  return 0;
}
