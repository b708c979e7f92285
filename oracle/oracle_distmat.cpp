// ORACLE / TEST INFRASTRUCTURE -- CPU restatement of the distance-matrix stage: RateModel::mlDistance (reference
// model.cpp:506-531), DistanceMatrixParams::tJC / tML (model.cpp:571-655) with GSL's golden-section minimiser
// (min/golden.c, as oracle/gslshim restates it) and distanceMatrixNegLogLike (model.cpp:555-569).  P_t comes from the
// model's eigen-decomposition (hist_model_desc), where the reference calls GSL's matrix exponential: the two agree to
// ~1e-13.  Pinned against the REAL reference's distanceMatrix on tests/golden/distmat_*.npz (tests/test_distmat.py): equal
// to 1e-6 relative wherever no comparison of the search ties, within its 1 % convergence tolerance otherwise.
// Only tests/ may use this file.
#include <cctype>
#include <cmath>
#include <complex>
#include <cstdint>
#include <map>
#include <vector>

#include "historian_b200.h"

namespace {
struct Params {
  const hist_model_desc* m;
  std::map<std::pair<int, int>, int> pairCount;
  double negLogLike(double t) const {   // model.cpp:555-569
    const int A = m->alph_size, C = m->components;
    double ll = 0;
    for (const auto& pc : pairCount) {
      const int i = pc.first.first, j = pc.first.second;
      double p = 0;
      for (int c = 0; c < C; ++c) {
        std::complex<double> acc = 0;
        for (int k = 0; k < A; ++k) {
          const size_t ik = ((size_t)c * A + i) * A + k, kj = ((size_t)c * A + k) * A + j, ck = (size_t)c * A + k;
          const std::complex<double> v(m->evec_re[ik], m->evec_im[ik]), vi(m->evec_inv_re[kj], m->evec_inv_im[kj]), lam(m->eval_re[ck], m->eval_im[ck]);
          acc += v * std::exp(lam * t) * vi;
        }
        p += m->cpt_weight[c] * acc.real();
      }
      ll += std::log(p) * (double)pc.second;
    }
    return -ll;
  }
};
}  // namespace

extern "C" int oracle_distance_matrix(const hist_model_desc* m, double rate, int32_t n, int32_t cols, const char* rows, int32_t maxIterations,
                                      double* dist) {
  const int A = m->alph_size;
  int lut[256];
  for (int c = 0; c < 256; ++c) lut[c] = -1;
  for (int k = 0; k < A; ++k) {   // tokenize, fastseq.cpp:10-16
    const unsigned char ch = (unsigned char)m->alphabet[k];
    const int other = isupper(ch) ? tolower(ch) : toupper(ch);
    if (lut[other] < 0) lut[other] = k;
  }
  for (int k = A; k-- > 0;) lut[(unsigned char)m->alphabet[k]] = k;
  lut['-'] = lut['.'] = lut['*'] = -1;
  for (int a = 0; a < n * n; ++a) dist[a] = 0;
  for (int x = 0; x + 1 < n; ++x)
    for (int y = x + 1; y < n; ++y) {
      Params dmp;
      dmp.m = m;
      for (int col = 0; col < cols; ++col) {   // mlDistance, model.cpp:506-520
        const int ti = lut[(unsigned char)rows[(size_t)x * cols + col]], tj = lut[(unsigned char)rows[(size_t)y * cols + col]];
        if (ti >= 0 && tj >= 0) ++dmp.pairCount[{ti, tj}];
      }
      // tJC, model.cpp:571-581
      int same = 0, diff = 0;
      for (const auto& pc : dmp.pairCount) (pc.first.first == pc.first.second ? same : diff) += pc.second;
      const double pDiff = diff / (double)(same + diff), dA = (double)A;
      const double tjcRaw = pDiff >= (dA - 1) / dA ? INFINITY : -((dA - 1) / dA) * std::log(1 - (dA / (dA - 1)) * pDiff) / rate;
      // tML, model.cpp:583-655
      const double tMin = 1e-9, tMax = 10;
      const double tjc = std::min(tMax, std::max(tMin, tjcRaw));
      double t = tjc;
      bool done = maxIterations <= 0;
      if (!done) {
        const double tLower = std::min(tMin, tjc / 2), tUpper = std::max(tMax, tjc * 2);
        const double llLower = dmp.negLogLike(tLower), llUpper = dmp.negLogLike(tUpper), lljc = dmp.negLogLike(tjc);
        if (!(lljc < llLower && lljc < llUpper)) {
          bool found = false;
          double lo = tLower, hi = tUpper;
          while (!found && hi - lo > tLower) {
            const double step = (hi - lo) / 4;
            for (double v = lo; v < hi && !found; v += step) {
              const double ll = dmp.negLogLike(v);
              if (ll < llLower && ll < llUpper) { found = true; t = v; }
            }
            if (!found) { if (llLower < llUpper) hi = (lo + hi) / 2; else lo = (lo + hi) / 2; }
          }
          if (!found) { t = llLower < llUpper ? tLower : tUpper; done = true; }
        }
        if (!done) {
          double xMin = t, xLo = tLower, xHi = tUpper, fMin = dmp.negLogLike(xMin);
          for (int iter = 0; iter < maxIterations; ++iter) {   // gsl_min_fminimizer_iterate, golden section
            const double wLower = xMin - xLo, wUpper = xHi - xMin;
            const double xNew = xMin + 0.3819660 * (wUpper > wLower ? wUpper : -wLower), fNew = dmp.negLogLike(xNew);
            if (fNew < fMin) { xMin = xNew; fMin = fNew; }
            else if (xNew < xMin && fNew > fMin) xLo = xNew;
            else if (xNew > xMin && fNew > fMin) xHi = xNew;
            t = xMin;
            const double minAbs = (xLo > 0 && xHi > 0) || (xLo < 0 && xHi < 0) ? std::min(std::fabs(xLo), std::fabs(xHi)) : 0;
            if (std::fabs(xHi - xLo) < 0.01 * minAbs) break;   // gsl_min_test_interval (a, b, 0, .01)
          }
        }
      }
      dist[(size_t)x * n + y] = dist[(size_t)y * n + x] = t;
    }
  return 0;
}
