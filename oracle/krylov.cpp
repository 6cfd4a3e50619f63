/* TEST INFRASTRUCTURE ONLY -- see krylov.h. */
#include "krylov.h"
#include <cmath>
#include <cstring>
#include <algorithm>

namespace nsoracle {

void Csr::mult(const double* x, double* y) const {
  for (int i = 0; i < n; i++) {
    double s = 0.0;
    for (int k = rowptr[i]; k < rowptr[i + 1]; k++) s += val[k] * x[col[k]];
    y[i] = s;
  }
}

void Csr::find_diag() {
  diag.assign(n, -1);
  for (int i = 0; i < n; i++)
    for (int k = rowptr[i]; k < rowptr[i + 1]; k++)
      if (col[k] == i) diag[i] = k;
}

/* ILU(0), natural ordering, IKJ variant on the pattern of A (columns sorted per row). */
void Ilu0::factor(const Csr& A_) {
  A = &A_;
  const Csr& a = A_;
  lu = a.val;
  std::vector<int> pos(a.n, -1);
  for (int i = 0; i < a.n; i++) {
    for (int k = a.rowptr[i]; k < a.rowptr[i + 1]; k++) pos[a.col[k]] = k;
    for (int k = a.rowptr[i]; k < a.rowptr[i + 1]; k++) {
      int j = a.col[k];
      if (j >= i) break;
      double piv = lu[a.diag[j]];
      double l = lu[k] / piv;
      lu[k] = l;
      for (int kk = a.diag[j] + 1; kk < a.rowptr[j + 1]; kk++) {
        int p = pos[a.col[kk]];
        if (p >= 0) lu[p] -= l * lu[kk];
      }
    }
    for (int k = a.rowptr[i]; k < a.rowptr[i + 1]; k++) pos[a.col[k]] = -1;
  }
}

void Ilu0::solve(const double* b, double* x) const {
  const Csr& a = *A;
  for (int i = 0; i < a.n; i++) {
    double s = b[i];
    for (int k = a.rowptr[i]; k < a.diag[i]; k++) s -= lu[k] * x[a.col[k]];
    x[i] = s;
  }
  for (int i = a.n - 1; i >= 0; i--) {
    double s = x[i];
    for (int k = a.diag[i] + 1; k < a.rowptr[i + 1]; k++) s -= lu[k] * x[a.col[k]];
    x[i] = s / lu[a.diag[i]];
  }
}

static double dot(int n, const double* a, const double* b) {
  double s = 0.0;
  for (int i = 0; i < n; i++) s += a[i] * b[i];
  return s;
}
static void axpy(int n, double a, const double* x, double* y) {
  for (int i = 0; i < n; i++) y[i] += a * x[i];
}
static void remove_mean(int n, double* x) {
  double s = 0.0;
  for (int i = 0; i < n; i++) s += x[i];
  s /= n;
  for (int i = 0; i < n; i++) x[i] -= s;
}

/* y = M^-1 A x (left-preconditioned operator), optionally with the constant projected out. */
struct PrecOp {
  const Csr& A;
  const Ilu0& M;
  bool nsp;
  std::vector<double> tmp;
  int matvecs = 0;
  PrecOp(const Csr& A_, const Ilu0& M_, bool nsp_) : A(A_), M(M_), nsp(nsp_), tmp(A_.n) {}
  void apply(const double* x, double* y) {
    A.mult(x, tmp.data());
    M.solve(tmp.data(), y);
    if (nsp) remove_mean(A.n, y);
    matvecs++;
  }
  void prec(const double* b, double* y) {
    M.solve(b, y);
    if (nsp) remove_mean(A.n, y);
  }
  /* r = M^-1 (b - A x) */
  void residual(const double* b, const double* x, double* r) {
    A.mult(x, tmp.data());
    for (int i = 0; i < A.n; i++) tmp[i] = b[i] - tmp[i];
    M.solve(tmp.data(), r);
    if (nsp) remove_mean(A.n, r);
    matvecs++;
  }
};

SolveStats gmres_ilu(const Csr& A, const Ilu0& M, const double* b, double* x, const SolveOpts& o) {
  const int n = A.n, m = o.restart;
  SolveStats st;
  PrecOp op(A, M, o.const_nullspace);
  std::vector<double> r(n), w(n);
  std::vector<std::vector<double>> V(m + 1, std::vector<double>(n));
  std::vector<double> H((m + 1) * m, 0.0), cs(m), sn(m), g(m + 1), hcol(m + 1);

  op.prec(b, r.data());
  st.rnorm0 = std::sqrt(dot(n, r.data(), r.data()));
  if (!o.nonzero_guess) std::memset(x, 0, sizeof(double) * n);
  else op.residual(b, x, r.data());
  const double target = std::max(o.rtol * st.rnorm0, o.abstol);
  double res = std::sqrt(dot(n, r.data(), r.data()));
  st.rnorm = res;
  if (res <= target) { st.converged = 1; st.matvecs = op.matvecs; return st; }

  while (st.iterations < o.maxits) {
    double beta = res;
    for (int i = 0; i < n; i++) V[0][i] = r[i] / beta;
    std::fill(g.begin(), g.end(), 0.0);
    g[0] = beta;
    int k = 0;
    for (; k < m && st.iterations < o.maxits; k++) {
      op.apply(V[k].data(), w.data());
      /* classical Gram-Schmidt: all projections from the same w */
      for (int j = 0; j <= k; j++) hcol[j] = dot(n, w.data(), V[j].data());
      for (int j = 0; j <= k; j++) axpy(n, -hcol[j], V[j].data(), w.data());
      double hn = std::sqrt(dot(n, w.data(), w.data()));
      hcol[k + 1] = hn;
      if (hn > 0) for (int i = 0; i < n; i++) V[k + 1][i] = w[i] / hn;
      for (int j = 0; j < k; j++) {
        double t = cs[j] * hcol[j] + sn[j] * hcol[j + 1];
        hcol[j + 1] = -sn[j] * hcol[j] + cs[j] * hcol[j + 1];
        hcol[j] = t;
      }
      double d = std::hypot(hcol[k], hcol[k + 1]);
      cs[k] = hcol[k] / d;
      sn[k] = hcol[k + 1] / d;
      hcol[k] = d;
      hcol[k + 1] = 0.0;
      g[k + 1] = -sn[k] * g[k];
      g[k] = cs[k] * g[k];
      for (int j = 0; j <= k; j++) H[j * m + k] = hcol[j];
      st.iterations++;
      res = std::fabs(g[k + 1]);
      if (res <= target || hn == 0.0) { k++; break; }
    }
    /* back substitution and update */
    std::vector<double> y(k);
    for (int i = k - 1; i >= 0; i--) {
      double s = g[i];
      for (int j = i + 1; j < k; j++) s -= H[i * m + j] * y[j];
      y[i] = s / H[i * m + i];
    }
    for (int j = 0; j < k; j++) axpy(n, y[j], V[j].data(), x);
    if (res <= target) { st.converged = 1; break; }
    op.residual(b, x, r.data());
    res = std::sqrt(dot(n, r.data(), r.data()));
    if (res <= target) { st.converged = 1; break; }
    if (res > o.dtol * st.rnorm0) break;
  }
  st.rnorm = res;
  st.matvecs = op.matvecs;
  return st;
}

/* BiCGStab(l), Sleijpen & Fokkema (ETNA 1, 1993), left-preconditioned. */
SolveStats bcgsl_ilu(const Csr& A, const Ilu0& M, const double* b, double* x, const SolveOpts& o) {
  const int n = A.n, L = o.ell;
  SolveStats st;
  PrecOp op(A, M, o.const_nullspace);
  std::vector<std::vector<double>> R(L + 1, std::vector<double>(n)), U(L + 1, std::vector<double>(n, 0.0));
  std::vector<double> rt(n);

  op.prec(b, R[0].data());
  st.rnorm0 = std::sqrt(dot(n, R[0].data(), R[0].data()));
  if (!o.nonzero_guess) std::memset(x, 0, sizeof(double) * n);
  else op.residual(b, x, R[0].data());
  const double target = std::max(o.rtol * st.rnorm0, o.abstol);
  double res = std::sqrt(dot(n, R[0].data(), R[0].data()));
  st.rnorm = res;
  if (res <= target) { st.converged = 1; st.matvecs = op.matvecs; return st; }
  rt = R[0];

  double rho0 = 1.0, alpha = 0.0, omega = 1.0;
  std::vector<double> tau((L + 1) * (L + 1), 0.0), sigma(L + 1), gam(L + 1), gamp(L + 1), gampp(L + 1);
  while (st.iterations < o.maxits) {
    rho0 = -omega * rho0;
    bool breakdown = false;
    for (int j = 0; j < L; j++) {
      double rho1 = dot(n, R[j].data(), rt.data());
      if (rho0 == 0.0) { breakdown = true; break; }
      double beta = alpha * rho1 / rho0;
      rho0 = rho1;
      for (int i = 0; i <= j; i++)
        for (int q = 0; q < n; q++) U[i][q] = R[i][q] - beta * U[i][q];
      op.apply(U[j].data(), U[j + 1].data());
      double gamma = dot(n, U[j + 1].data(), rt.data());
      if (gamma == 0.0) { breakdown = true; break; }
      alpha = rho0 / gamma;
      for (int i = 0; i <= j; i++) axpy(n, -alpha, U[i + 1].data(), R[i].data());
      op.apply(R[j].data(), R[j + 1].data());
      axpy(n, alpha, U[0].data(), x);
      st.iterations++;
    }
    if (breakdown) break;
    for (int j = 1; j <= L; j++) {
      for (int i = 1; i < j; i++) {
        double t = dot(n, R[j].data(), R[i].data()) / sigma[i];
        tau[i * (L + 1) + j] = t;
        axpy(n, -t, R[i].data(), R[j].data());
      }
      sigma[j] = dot(n, R[j].data(), R[j].data());
      if (sigma[j] == 0.0) { breakdown = true; break; }
      gamp[j] = dot(n, R[0].data(), R[j].data()) / sigma[j];
    }
    if (breakdown) break;
    gam[L] = gamp[L];
    omega = gam[L];
    for (int j = L - 1; j >= 1; j--) {
      double s = gamp[j];
      for (int i = j + 1; i <= L; i++) s -= tau[j * (L + 1) + i] * gam[i];
      gam[j] = s;
    }
    for (int j = 1; j < L; j++) {
      double s = gam[j + 1];
      for (int i = j + 1; i < L; i++) s += tau[j * (L + 1) + i] * gam[i + 1];
      gampp[j] = s;
    }
    axpy(n, gam[1], R[0].data(), x);
    axpy(n, -gamp[L], R[L].data(), R[0].data());
    axpy(n, -gam[L], U[L].data(), U[0].data());
    for (int j = 1; j < L; j++) {
      axpy(n, -gam[j], U[j].data(), U[0].data());
      axpy(n, gampp[j], R[j].data(), x);
      axpy(n, -gamp[j], R[j].data(), R[0].data());
    }
    res = std::sqrt(dot(n, R[0].data(), R[0].data()));
    if (res <= target) { st.converged = 1; break; }
    if (res > o.dtol * st.rnorm0 || !(res == res)) break;
  }
  st.rnorm = res;
  st.matvecs = op.matvecs;
  return st;
}

}  // namespace nsoracle
