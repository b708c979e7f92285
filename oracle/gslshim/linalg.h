// Small dense linear algebra for the host side (no GSL, no LAPACK).
//
// The reference leans on GSL for five numerical routines on the hot path's
// model-preparation side: matrix exponential (model.cpp:322-334), nonsymmetric
// eigendecomposition + complex LU inverse (model.cpp:1155-1170), QR least
// squares (model.cpp:282-320) and the gamma quantile (gamma.cpp:15).  This
// header provides from-scratch equivalents on plain row-major arrays so the
// host (and the oracle's GSL shim, which wraps these) needs no external library.
#pragma once
#include <algorithm>
#include <cmath>
#include <complex>
#include <limits>
#include <vector>

namespace hb {
namespace linalg {

typedef std::complex<double> cplx;

// ---------------------------------------------------------------- expm
// exp(A) for an n x n row-major real matrix: scaling and squaring around a
// degree-18 Taylor polynomial evaluated at ||A/2^s||_1 <= 1/8.
inline void expm(int n, const double* A, double* out) {
  double norm1 = 0;
  for (int j = 0; j < n; ++j) {
    double c = 0;
    for (int i = 0; i < n; ++i) c += std::fabs(A[i * n + j]);
    norm1 = std::max(norm1, c);
  }
  int s = 0;
  if (norm1 > 0.125) s = std::max(0, (int)std::ceil(std::log2(norm1 / 0.125)));
  const double scale = std::ldexp(1.0, -s);
  std::vector<double> B(n * n), term(n * n), sum(n * n), tmp(n * n);
  for (int i = 0; i < n * n; ++i) B[i] = A[i] * scale;
  for (int i = 0; i < n * n; ++i) term[i] = sum[i] = 0;
  for (int i = 0; i < n; ++i) term[i * n + i] = sum[i * n + i] = 1;
  for (int k = 1; k <= 18; ++k) {
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) {
        double acc = 0;
        for (int m = 0; m < n; ++m) acc += term[i * n + m] * B[m * n + j];
        tmp[i * n + j] = acc / k;
      }
    term.swap(tmp);
    for (int i = 0; i < n * n; ++i) sum[i] += term[i];
  }
  for (int q = 0; q < s; ++q) {
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) {
        double acc = 0;
        for (int m = 0; m < n; ++m) acc += sum[i * n + m] * sum[m * n + j];
        tmp[i * n + j] = acc;
      }
    sum.swap(tmp);
  }
  for (int i = 0; i < n * n; ++i) out[i] = sum[i];
}

// ---------------------------------------------------------------- eigen
// Eigen-decomposition of a real nonsymmetric n x n matrix (row-major):
// A V = V diag(w).  Complex Hessenberg reduction, single-shift QR to Schur
// form, triangular back-substitution, unit 2-norm eigenvectors (columns of V,
// V[i*n+k] = component i of eigenvector k).  Returns false if QR fails to
// converge.
inline bool eig_nonsymm(int n, const double* A, cplx* w, cplx* V) {
  std::vector<cplx> H(n * n), Q(n * n, cplx(0, 0));
  for (int i = 0; i < n * n; ++i) H[i] = A[i];
  for (int i = 0; i < n; ++i) Q[i * n + i] = 1;
  const double eps = std::numeric_limits<double>::epsilon();
  // Householder reduction to upper Hessenberg form
  std::vector<cplx> v(n);
  for (int k = 0; k + 2 < n; ++k) {
    double xn = 0;
    for (int i = k + 1; i < n; ++i) xn += std::norm(H[i * n + k]);
    xn = std::sqrt(xn);
    if (xn == 0) continue;
    const cplx x0 = H[(k + 1) * n + k];
    const cplx phase = std::abs(x0) > 0 ? x0 / std::abs(x0) : cplx(1, 0);
    const cplx alpha = -phase * xn;
    double vn = 0;
    for (int i = k + 1; i < n; ++i) {
      v[i] = H[i * n + k] - (i == k + 1 ? alpha : cplx(0, 0));
      vn += std::norm(v[i]);
    }
    if (vn == 0) continue;
    vn = std::sqrt(vn);
    for (int i = k + 1; i < n; ++i) v[i] /= vn;
    // H <- P H, P = I - 2 v v^H
    for (int j = 0; j < n; ++j) {
      cplx d = 0;
      for (int i = k + 1; i < n; ++i) d += std::conj(v[i]) * H[i * n + j];
      d *= 2.0;
      for (int i = k + 1; i < n; ++i) H[i * n + j] -= v[i] * d;
    }
    // H <- H P ; Q <- Q P
    for (int i = 0; i < n; ++i) {
      cplx d = 0, e = 0;
      for (int j = k + 1; j < n; ++j) {
        d += H[i * n + j] * v[j];
        e += Q[i * n + j] * v[j];
      }
      d *= 2.0;
      e *= 2.0;
      for (int j = k + 1; j < n; ++j) {
        H[i * n + j] -= d * std::conj(v[j]);
        Q[i * n + j] -= e * std::conj(v[j]);
      }
    }
    for (int i = k + 2; i < n; ++i) H[i * n + k] = 0;
  }
  // shifted QR on the active window [l, hi]
  std::vector<cplx> gc(n), gs(n);
  int hi = n - 1, iter = 0, total = 0;
  while (hi > 0) {
    int l = hi;
    while (l > 0) {
      const double sub = std::abs(H[l * n + l - 1]);
      double diag = std::abs(H[(l - 1) * n + l - 1]) + std::abs(H[l * n + l]);
      if (diag == 0) diag = 1;
      if (sub <= eps * diag) {
        H[l * n + l - 1] = 0;
        break;
      }
      --l;
    }
    if (l == hi) {
      --hi;
      iter = 0;
      continue;
    }
    if (++total > 300 * n) return false;
    ++iter;
    // Wilkinson shift: eigenvalue of the trailing 2x2 closest to H[hi][hi]
    const cplx a = H[(hi - 1) * n + hi - 1], b = H[(hi - 1) * n + hi];
    const cplx c = H[hi * n + hi - 1], d = H[hi * n + hi];
    cplx mu;
    if (iter % 11 == 10) {
      mu = d + cplx(std::abs(c), 0.0);  // exceptional shift
    } else {
      const cplx tr = a + d, det = a * d - b * c;
      const cplx disc = std::sqrt(tr * tr * 0.25 - det);
      const cplx m1 = tr * 0.5 + disc, m2 = tr * 0.5 - disc;
      mu = std::abs(m1 - d) < std::abs(m2 - d) ? m1 : m2;
    }
    for (int i = l; i <= hi; ++i) H[i * n + i] -= mu;
    for (int k = l; k < hi; ++k) {
      const cplx p = H[k * n + k], q = H[(k + 1) * n + k];
      const double r = std::sqrt(std::norm(p) + std::norm(q));
      if (r == 0) {
        gc[k] = 1;
        gs[k] = 0;
        continue;
      }
      const cplx cc = p / r, ss = q / r;
      gc[k] = cc;
      gs[k] = ss;
      for (int j = k; j < n; ++j) {
        const cplx t1 = H[k * n + j], t2 = H[(k + 1) * n + j];
        H[k * n + j] = std::conj(cc) * t1 + std::conj(ss) * t2;
        H[(k + 1) * n + j] = -ss * t1 + cc * t2;
      }
      H[(k + 1) * n + k] = 0;
    }
    for (int k = l; k < hi; ++k) {
      const cplx cc = gc[k], ss = gs[k];
      const int rmax = std::min(k + 2, hi);
      for (int i = 0; i <= rmax; ++i) {
        const cplx t1 = H[i * n + k], t2 = H[i * n + k + 1];
        H[i * n + k] = t1 * cc + t2 * ss;
        H[i * n + k + 1] = -std::conj(ss) * t1 + std::conj(cc) * t2;
      }
      for (int i = 0; i < n; ++i) {
        const cplx t1 = Q[i * n + k], t2 = Q[i * n + k + 1];
        Q[i * n + k] = t1 * cc + t2 * ss;
        Q[i * n + k + 1] = -std::conj(ss) * t1 + std::conj(cc) * t2;
      }
    }
    for (int i = l; i <= hi; ++i) H[i * n + i] += mu;
  }
  // eigenvectors of the triangular factor, then rotate back by Q
  double tnorm = 0;
  for (int i = 0; i < n; ++i)
    for (int j = i; j < n; ++j) tnorm = std::max(tnorm, std::abs(H[i * n + j]));
  const double smin = std::max(eps * tnorm, std::numeric_limits<double>::min());
  std::vector<cplx> y(n);
  for (int k = 0; k < n; ++k) {
    const cplx lam = H[k * n + k];
    w[k] = lam;
    for (int j = 0; j < n; ++j) y[j] = 0;
    y[k] = 1;
    for (int j = k - 1; j >= 0; --j) {
      cplx acc = 0;
      for (int m = j + 1; m <= k; ++m) acc += H[j * n + m] * y[m];
      cplx den = H[j * n + j] - lam;
      if (std::abs(den) < smin) den = smin;
      y[j] = -acc / den;
    }
    double nrm = 0;
    for (int i = 0; i < n; ++i) {
      cplx acc = 0;
      for (int j = 0; j <= k; ++j) acc += Q[i * n + j] * y[j];
      V[i * n + k] = acc;
      nrm += std::norm(acc);
    }
    nrm = std::sqrt(nrm);
    if (nrm > 0)
      for (int i = 0; i < n; ++i) V[i * n + k] /= nrm;
  }
  // a real input has real eigenvectors for real eigenvalues: strip the
  // arbitrary complex phase so that callers testing "is this model real?" see it
  for (int k = 0; k < n; ++k) {
    if (std::fabs(w[k].imag()) > 1e-12 * std::max(1.0, std::abs(w[k]))) continue;
    int big = 0;
    for (int i = 1; i < n; ++i)
      if (std::abs(V[i * n + k]) > std::abs(V[big * n + k])) big = i;
    const cplx ph = V[big * n + k];
    if (std::abs(ph) == 0) continue;
    const cplx rot = std::conj(ph) / std::abs(ph);
    for (int i = 0; i < n; ++i) V[i * n + k] *= rot;
  }
  return true;
}

// ---------------------------------------------------------------- complex inverse
// Inverse of an n x n complex matrix by Gauss-Jordan with partial pivoting.
inline bool cinverse(int n, const cplx* A, cplx* Ainv) {
  std::vector<cplx> M(A, A + n * n);
  for (int i = 0; i < n * n; ++i) Ainv[i] = 0;
  for (int i = 0; i < n; ++i) Ainv[i * n + i] = 1;
  for (int c = 0; c < n; ++c) {
    int p = c;
    for (int r = c + 1; r < n; ++r)
      if (std::abs(M[r * n + c]) > std::abs(M[p * n + c])) p = r;
    if (std::abs(M[p * n + c]) == 0) return false;
    if (p != c)
      for (int j = 0; j < n; ++j) {
        std::swap(M[p * n + j], M[c * n + j]);
        std::swap(Ainv[p * n + j], Ainv[c * n + j]);
      }
    const cplx piv = M[c * n + c];
    for (int j = 0; j < n; ++j) {
      M[c * n + j] /= piv;
      Ainv[c * n + j] /= piv;
    }
    for (int r = 0; r < n; ++r) {
      if (r == c) continue;
      const cplx f = M[r * n + c];
      if (f == cplx(0, 0)) continue;
      for (int j = 0; j < n; ++j) {
        M[r * n + j] -= f * M[c * n + j];
        Ainv[r * n + j] -= f * Ainv[c * n + j];
      }
    }
  }
  return true;
}

// ---------------------------------------------------------------- QR least squares
// Householder QR of an m x n (m >= n) row-major matrix in place: R in the upper
// triangle, Householder vectors (v[0] = 1 implicit) below it, scalars in tau.
inline void qr_decomp(int m, int n, double* A, double* tau) {
  for (int k = 0; k < n; ++k) {
    double xn = 0;
    for (int i = k; i < m; ++i) xn += A[i * n + k] * A[i * n + k];
    xn = std::sqrt(xn);
    if (xn == 0) {
      tau[k] = 0;
      continue;
    }
    const double alpha = A[k * n + k] > 0 ? -xn : xn;
    const double v0 = A[k * n + k] - alpha;
    tau[k] = -v0 / alpha;  // = 2 / (v^T v) with v scaled so v[0] = 1
    for (int i = k + 1; i < m; ++i) A[i * n + k] /= v0;
    A[k * n + k] = alpha;
    for (int j = k + 1; j < n; ++j) {
      double d = A[k * n + j];
      for (int i = k + 1; i < m; ++i) d += A[i * n + k] * A[i * n + j];
      d *= tau[k];
      A[k * n + j] -= d;
      for (int i = k + 1; i < m; ++i) A[i * n + j] -= d * A[i * n + k];
    }
  }
}

// Least-squares solve min ||A x - b|| from qr_decomp output; residual = b - A x.
inline void qr_lssolve(int m, int n, const double* QR, const double* tau, const double* b,
                       double* x, double* residual) {
  std::vector<double> c(b, b + m);
  for (int k = 0; k < n; ++k) {  // c <- Q^T b
    double d = c[k];
    for (int i = k + 1; i < m; ++i) d += QR[i * n + k] * c[i];
    d *= tau[k];
    c[k] -= d;
    for (int i = k + 1; i < m; ++i) c[i] -= d * QR[i * n + k];
  }
  for (int k = n - 1; k >= 0; --k) {
    double acc = c[k];
    for (int j = k + 1; j < n; ++j) acc -= QR[k * n + j] * x[j];
    x[k] = acc / QR[k * n + k];
  }
  if (residual) {
    std::vector<double> r(m, 0.0);
    for (int i = n; i < m; ++i) r[i] = c[i];
    for (int k = n - 1; k >= 0; --k) {  // r <- Q r
      double d = r[k];
      for (int i = k + 1; i < m; ++i) d += QR[i * n + k] * r[i];
      d *= tau[k];
      r[k] -= d;
      for (int i = k + 1; i < m; ++i) r[i] -= d * QR[i * n + k];
    }
    for (int i = 0; i < m; ++i) residual[i] = r[i];
  }
}

// ---------------------------------------------------------------- gamma functions
// Regularised lower incomplete gamma P(a, x).
inline double gamma_p(double a, double x) {
  if (x <= 0) return 0;
  const double gln = std::lgamma(a);
  if (x < a + 1) {
    double ap = a, sum = 1.0 / a, del = sum;
    for (int n = 0; n < 10000; ++n) {
      ap += 1;
      del *= x / ap;
      sum += del;
      if (std::fabs(del) < std::fabs(sum) * 1e-17) break;
    }
    return sum * std::exp(-x + a * std::log(x) - gln);
  }
  const double tiny = 1e-300;
  double b = x + 1 - a, c = 1 / tiny, d = 1 / b, h = d;
  for (int i = 1; i < 10000; ++i) {
    const double an = -i * (i - a);
    b += 2;
    d = an * d + b;
    if (std::fabs(d) < tiny) d = tiny;
    c = b + an / c;
    if (std::fabs(c) < tiny) c = tiny;
    d = 1 / d;
    const double del = d * c;
    h *= del;
    if (std::fabs(del - 1) < 1e-16) break;
  }
  return 1 - std::exp(-x + a * std::log(x) - gln) * h;
}

// Quantile of Gamma(shape a, scale b): x with P(a, x / b) = p.
inline double gamma_pinv(double p, double a, double b) {
  if (p <= 0) return 0;
  if (p >= 1) return std::numeric_limits<double>::infinity();
  double lo = 0, hi = std::max(1.0, a);
  while (gamma_p(a, hi) < p) hi *= 2;
  for (int it = 0; it < 200; ++it) {
    const double mid = 0.5 * (lo + hi);
    if (gamma_p(a, mid) < p) lo = mid; else hi = mid;
    if (hi - lo <= 1e-16 * hi) break;
  }
  double x = 0.5 * (lo + hi);
  for (int it = 0; it < 4; ++it) {  // Newton polish
    const double f = gamma_p(a, x) - p;
    const double dens = std::exp((a - 1) * std::log(x) - x - std::lgamma(a));
    if (!(dens > 0)) break;
    const double xn = x - f / dens;
    if (!(xn > lo && xn < hi)) break;
    x = xn;
  }
  return x * b;
}

}  // namespace linalg
}  // namespace hb
