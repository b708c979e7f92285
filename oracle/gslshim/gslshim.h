// GSL shim — ORACLE / TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// GSL is absent from this image, and it is the only thing between us and a
// compiled copy of the unmodified reference.  This header
// supplies exactly the GSL types and functions the reference's sources touch
// (containers, complex helpers, expm, nonsymmetric eigensolver, complex LU
// inverse, QR least squares, golden-section minimiser, gamma quantile, three
// pdfs, gsl_fcmp), header-only, on top of oracle/gslshim/linalg.h (dense linear algebra written for this shim).
// The reference is then compiled from /root/reference/src where it lies by
// oracle/Makefile into oracle/_ref/.
#pragma once
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <complex>
#include <vector>
#include "linalg.h"

#define GSL_SUCCESS 0
#define GSL_FAILURE (-1)
#define GSL_CONTINUE (-2)
#define GSL_EINVAL 4
#define GSL_PREC_DOUBLE 0
typedef unsigned int gsl_mode_t;

// ---- containers -----------------------------------------------------------
struct gsl_vector { size_t size; double* data; };
struct gsl_matrix { size_t size1, size2; double* data; };
struct gsl_complex { double dat[2]; };
struct gsl_vector_complex { size_t size; gsl_complex* data; };
struct gsl_matrix_complex { size_t size1, size2; gsl_complex* data; };
struct gsl_permutation { size_t size; size_t* data; };

#define GSL_REAL(z) ((z).dat[0])
#define GSL_IMAG(z) ((z).dat[1])

inline gsl_vector* gsl_vector_alloc(size_t n) {
  gsl_vector* v = new gsl_vector;
  v->size = n;
  v->data = new double[n ? n : 1]();
  return v;
}
inline void gsl_vector_free(gsl_vector* v) { if (v) { delete[] v->data; delete v; } }
inline double gsl_vector_get(const gsl_vector* v, size_t i) { return v->data[i]; }
inline void gsl_vector_set(gsl_vector* v, size_t i, double x) { v->data[i] = x; }
inline void gsl_vector_set_zero(gsl_vector* v) { for (size_t i = 0; i < v->size; ++i) v->data[i] = 0; }
inline int gsl_vector_memcpy(gsl_vector* d, const gsl_vector* s) {
  if (d->size != s->size) return GSL_EINVAL;
  std::memcpy(d->data, s->data, s->size * sizeof(double));
  return GSL_SUCCESS;
}
inline int gsl_vector_scale(gsl_vector* v, double x) { for (size_t i = 0; i < v->size; ++i) v->data[i] *= x; return GSL_SUCCESS; }

inline gsl_matrix* gsl_matrix_alloc(size_t n1, size_t n2) {
  gsl_matrix* m = new gsl_matrix;
  m->size1 = n1;
  m->size2 = n2;
  m->data = new double[n1 * n2 ? n1 * n2 : 1]();
  return m;
}
inline void gsl_matrix_free(gsl_matrix* m) { if (m) { delete[] m->data; delete m; } }
inline double gsl_matrix_get(const gsl_matrix* m, size_t i, size_t j) { return m->data[i * m->size2 + j]; }
inline void gsl_matrix_set(gsl_matrix* m, size_t i, size_t j, double x) { m->data[i * m->size2 + j] = x; }
inline double* gsl_matrix_ptr(gsl_matrix* m, size_t i, size_t j) { return m->data + i * m->size2 + j; }
inline void gsl_matrix_set_zero(gsl_matrix* m) { for (size_t i = 0; i < m->size1 * m->size2; ++i) m->data[i] = 0; }
inline int gsl_matrix_memcpy(gsl_matrix* d, const gsl_matrix* s) {
  if (d->size1 != s->size1 || d->size2 != s->size2) return GSL_EINVAL;
  std::memcpy(d->data, s->data, s->size1 * s->size2 * sizeof(double));
  return GSL_SUCCESS;
}
inline int gsl_matrix_scale(gsl_matrix* m, double x) { for (size_t i = 0; i < m->size1 * m->size2; ++i) m->data[i] *= x; return GSL_SUCCESS; }

inline gsl_vector_complex* gsl_vector_complex_alloc(size_t n) {
  gsl_vector_complex* v = new gsl_vector_complex;
  v->size = n;
  v->data = new gsl_complex[n ? n : 1]();
  return v;
}
inline void gsl_vector_complex_free(gsl_vector_complex* v) { if (v) { delete[] v->data; delete v; } }
inline gsl_complex gsl_vector_complex_get(const gsl_vector_complex* v, size_t i) { return v->data[i]; }
inline void gsl_vector_complex_set(gsl_vector_complex* v, size_t i, gsl_complex z) { v->data[i] = z; }
inline int gsl_vector_complex_memcpy(gsl_vector_complex* d, const gsl_vector_complex* s) {
  if (d->size != s->size) return GSL_EINVAL;
  std::memcpy(d->data, s->data, s->size * sizeof(gsl_complex));
  return GSL_SUCCESS;
}

inline gsl_matrix_complex* gsl_matrix_complex_alloc(size_t n1, size_t n2) {
  gsl_matrix_complex* m = new gsl_matrix_complex;
  m->size1 = n1;
  m->size2 = n2;
  m->data = new gsl_complex[n1 * n2 ? n1 * n2 : 1]();
  return m;
}
inline void gsl_matrix_complex_free(gsl_matrix_complex* m) { if (m) { delete[] m->data; delete m; } }
inline gsl_complex gsl_matrix_complex_get(const gsl_matrix_complex* m, size_t i, size_t j) { return m->data[i * m->size2 + j]; }
inline void gsl_matrix_complex_set(gsl_matrix_complex* m, size_t i, size_t j, gsl_complex z) { m->data[i * m->size2 + j] = z; }
inline int gsl_matrix_complex_memcpy(gsl_matrix_complex* d, const gsl_matrix_complex* s) {
  if (d->size1 != s->size1 || d->size2 != s->size2) return GSL_EINVAL;
  std::memcpy(d->data, s->data, s->size1 * s->size2 * sizeof(gsl_complex));
  return GSL_SUCCESS;
}

inline gsl_permutation* gsl_permutation_alloc(size_t n) {
  gsl_permutation* p = new gsl_permutation;
  p->size = n;
  p->data = new size_t[n ? n : 1]();
  return p;
}
inline void gsl_permutation_free(gsl_permutation* p) { if (p) { delete[] p->data; delete p; } }

// ---- complex arithmetic -----------------------------------------------------
inline gsl_complex gsl_complex_rect(double x, double y) { gsl_complex z; z.dat[0] = x; z.dat[1] = y; return z; }
inline std::complex<double> gslshim_c(gsl_complex z) { return std::complex<double>(z.dat[0], z.dat[1]); }
inline gsl_complex gslshim_g(std::complex<double> z) { return gsl_complex_rect(z.real(), z.imag()); }
inline gsl_complex gsl_complex_add(gsl_complex a, gsl_complex b) { return gsl_complex_rect(a.dat[0] + b.dat[0], a.dat[1] + b.dat[1]); }
inline gsl_complex gsl_complex_sub(gsl_complex a, gsl_complex b) { return gsl_complex_rect(a.dat[0] - b.dat[0], a.dat[1] - b.dat[1]); }
inline gsl_complex gsl_complex_mul(gsl_complex a, gsl_complex b) {
  return gsl_complex_rect(a.dat[0] * b.dat[0] - a.dat[1] * b.dat[1], a.dat[0] * b.dat[1] + a.dat[1] * b.dat[0]);
}
inline gsl_complex gsl_complex_div(gsl_complex a, gsl_complex b) { return gslshim_g(gslshim_c(a) / gslshim_c(b)); }
inline gsl_complex gsl_complex_mul_real(gsl_complex a, double x) { return gsl_complex_rect(a.dat[0] * x, a.dat[1] * x); }
inline gsl_complex gsl_complex_exp(gsl_complex a) {
  const double rho = std::exp(a.dat[0]);
  return gsl_complex_rect(rho * std::cos(a.dat[1]), rho * std::sin(a.dat[1]));
}

// ---- linear algebra ---------------------------------------------------------
inline int gsl_linalg_exponential_ss(const gsl_matrix* A, gsl_matrix* eA, gsl_mode_t) {
  if (A->size1 != A->size2 || eA->size1 != A->size1 || eA->size2 != A->size2) return GSL_EINVAL;
  hb::linalg::expm((int)A->size1, A->data, eA->data);
  return GSL_SUCCESS;
}

struct gsl_eigen_nonsymmv_workspace { size_t size; };
inline gsl_eigen_nonsymmv_workspace* gsl_eigen_nonsymmv_alloc(size_t n) {
  gsl_eigen_nonsymmv_workspace* w = new gsl_eigen_nonsymmv_workspace;
  w->size = n;
  return w;
}
inline void gsl_eigen_nonsymmv_free(gsl_eigen_nonsymmv_workspace* w) { delete w; }
inline int gsl_eigen_nonsymmv(gsl_matrix* A, gsl_vector_complex* eval, gsl_matrix_complex* evec, gsl_eigen_nonsymmv_workspace*) {
  const int n = (int)A->size1;
  std::vector<hb::linalg::cplx> w(n), V(n * n);
  if (!hb::linalg::eig_nonsymm(n, A->data, w.data(), V.data())) return GSL_FAILURE;
  for (int i = 0; i < n; ++i) eval->data[i] = gslshim_g(w[i]);
  for (int i = 0; i < n * n; ++i) evec->data[i] = gslshim_g(V[i]);
  return GSL_SUCCESS;
}

// The reference only ever calls LU_decomp immediately followed by LU_invert
// (model.cpp:1167-1168), so the "decomposition" keeps the matrix and the
// inversion does the elimination.
inline int gsl_linalg_complex_LU_decomp(gsl_matrix_complex*, gsl_permutation*, int* signum) {
  if (signum) *signum = 1;
  return GSL_SUCCESS;
}
inline int gsl_linalg_complex_LU_invert(const gsl_matrix_complex* LU, const gsl_permutation*, gsl_matrix_complex* inverse) {
  const int n = (int)LU->size1;
  std::vector<hb::linalg::cplx> a(n * n), ai(n * n);
  for (int i = 0; i < n * n; ++i) a[i] = gslshim_c(LU->data[i]);
  if (!hb::linalg::cinverse(n, a.data(), ai.data())) return GSL_FAILURE;
  for (int i = 0; i < n * n; ++i) inverse->data[i] = gslshim_g(ai[i]);
  return GSL_SUCCESS;
}

inline int gsl_linalg_QR_decomp(gsl_matrix* A, gsl_vector* tau) {
  hb::linalg::qr_decomp((int)A->size1, (int)A->size2, A->data, tau->data);
  return GSL_SUCCESS;
}
inline int gsl_linalg_QR_lssolve(const gsl_matrix* QR, const gsl_vector* tau, const gsl_vector* b, gsl_vector* x, gsl_vector* residual) {
  hb::linalg::qr_lssolve((int)QR->size1, (int)QR->size2, QR->data, tau->data, b->data, x->data, residual ? residual->data : 0);
  return GSL_SUCCESS;
}

// ---- 1-D minimisation (golden section, GSL semantics) -------------------------
struct gsl_function { double (*function)(double x, void* params); void* params; };
struct gsl_min_fminimizer_type { int kind; };
static const gsl_min_fminimizer_type gslshim_golden_type = {0};
static const gsl_min_fminimizer_type gslshim_brent_type = {1};
static const gsl_min_fminimizer_type* const gsl_min_fminimizer_goldensection = &gslshim_golden_type;
static const gsl_min_fminimizer_type* const gsl_min_fminimizer_brent = &gslshim_brent_type;
struct gsl_min_fminimizer {
  const gsl_min_fminimizer_type* type;
  gsl_function* f;
  double x_minimum, x_lower, x_upper, f_minimum, f_lower, f_upper;
};
inline gsl_min_fminimizer* gsl_min_fminimizer_alloc(const gsl_min_fminimizer_type* T) {
  gsl_min_fminimizer* s = new gsl_min_fminimizer();
  s->type = T;
  return s;
}
inline void gsl_min_fminimizer_free(gsl_min_fminimizer* s) { delete s; }
inline int gsl_min_fminimizer_set(gsl_min_fminimizer* s, gsl_function* f, double x_minimum, double x_lower, double x_upper) {
  s->f = f;
  s->x_minimum = x_minimum;
  s->x_lower = x_lower;
  s->x_upper = x_upper;
  s->f_minimum = f->function(x_minimum, f->params);
  s->f_lower = f->function(x_lower, f->params);
  s->f_upper = f->function(x_upper, f->params);
  if (s->f_minimum >= s->f_lower || s->f_minimum >= s->f_upper) return GSL_EINVAL;
  return GSL_SUCCESS;
}
inline int gsl_min_fminimizer_iterate(gsl_min_fminimizer* s) {
  const double x_center = s->x_minimum, x_left = s->x_lower, x_right = s->x_upper;
  const double f_min = s->f_minimum;
  const double golden = 0.3819660;
  const double w_lower = x_center - x_left, w_upper = x_right - x_center;
  const double x_new = x_center + golden * ((w_upper > w_lower) ? w_upper : -w_lower);
  const double f_new = s->f->function(x_new, s->f->params);
  if (f_new < f_min) {
    s->x_minimum = x_new;
    s->f_minimum = f_new;
    return GSL_SUCCESS;
  } else if (x_new < x_center && f_new > f_min) {
    s->x_lower = x_new;
    s->f_lower = f_new;
    return GSL_SUCCESS;
  } else if (x_new > x_center && f_new > f_min) {
    s->x_upper = x_new;
    s->f_upper = f_new;
    return GSL_SUCCESS;
  }
  return GSL_FAILURE;
}
inline double gsl_min_fminimizer_x_minimum(const gsl_min_fminimizer* s) { return s->x_minimum; }
inline double gsl_min_fminimizer_x_lower(const gsl_min_fminimizer* s) { return s->x_lower; }
inline double gsl_min_fminimizer_x_upper(const gsl_min_fminimizer* s) { return s->x_upper; }
inline int gsl_min_test_interval(double x_lower, double x_upper, double epsabs, double epsrel) {
  const double abs_lower = std::fabs(x_lower), abs_upper = std::fabs(x_upper);
  double min_abs = 0;
  if ((x_lower > 0 && x_upper > 0) || (x_lower < 0 && x_upper < 0)) min_abs = abs_lower < abs_upper ? abs_lower : abs_upper;
  const double tolerance = epsabs + epsrel * min_abs;
  return std::fabs(x_upper - x_lower) < tolerance ? GSL_SUCCESS : GSL_CONTINUE;
}

// ---- distributions ------------------------------------------------------------
inline double gsl_cdf_gamma_Pinv(double P, double a, double b) { return hb::linalg::gamma_pinv(P, a, b); }
inline double gsl_ran_beta_pdf(double x, double a, double b) {
  if (x < 0 || x > 1) return 0;
  return std::exp(std::lgamma(a + b) - std::lgamma(a) - std::lgamma(b) + (a - 1) * std::log(x) + (b - 1) * std::log1p(-x));
}
inline double gsl_ran_gamma_pdf(double x, double a, double b) {
  if (x < 0) return 0;
  if (x == 0) return a == 1 ? 1 / b : 0;
  return std::exp((a - 1) * std::log(x / b) - x / b - std::lgamma(a)) / b;
}
inline double gsl_ran_dirichlet_pdf(size_t K, const double alpha[], const double theta[]) {
  double lp = 0, sum = 0;
  for (size_t i = 0; i < K; ++i) {
    lp += (alpha[i] - 1) * std::log(theta[i]) - std::lgamma(alpha[i]);
    sum += alpha[i];
  }
  return std::exp(lp + std::lgamma(sum));
}

// ---- misc ------------------------------------------------------------------------
inline int gsl_fcmp(double x1, double x2, double epsilon) {
  int exponent;
  const double mx = (std::fabs(x1) > std::fabs(x2)) ? x1 : x2;
  std::frexp(mx, &exponent);
  const double delta = std::ldexp(epsilon, exponent);
  const double difference = x1 - x2;
  if (difference > delta) return 1;
  if (difference < -delta) return -1;
  return 0;
}
inline const char* gsl_strerror(int code) { return code == 0 ? "success" : "gsl shim error"; }
