// TEST INFRASTRUCTURE (CPU oracle) — not part of the product path.
// Preconditioned conjugate gradients as lib/small_thing/PCG.cpp:15-97 runs them, and the block-Jacobi preconditioner of
// lib/small_thing/Preconditioner.h:52-116, restated over plain vectors so that the same routine serves the oracle's
// reduced camera system (solver.hpp) and the dense systems of the reference's own PCG test recipe
// (lib/small_thing/tests/TestPCG.cpp:29-131, restated in tests/test_pcg.py).
#pragma once
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <vector>

namespace orc {

struct PcgResult { int numIterations; double relativeResidual; };

// PCG::solve: x_0 = 0, r_0 = b, z_0 = M^-1 r_0, p_0 = z_0; per step alpha = r.z / p.Ap, x += alpha p, r -= alpha Ap,
// exit on |r| / |r_0| < desired (PCG.cpp:60-62) or after maxIterations products (:63-69), then z' = M^-1 r,
// beta = z'.r' / z.r, p = z' + beta p.
template <typename ApplyA, typename ApplyInvM>
PcgResult pcgRun(ApplyA&& applyA, ApplyInvM&& applyInvM, const std::vector<double>& b, double desired, int maxIterations,
                 std::vector<double>& x) {
  const int64_t n = (int64_t)b.size();
  std::vector<double> xk((size_t)n, 0.0), rk(b), zk, pk, Apk, zk1;
  auto dot = [&](const std::vector<double>& a, const std::vector<double>& c) { double s = 0; for (int64_t i = 0; i < n; i++) s += a[i] * c[i]; return s; };
  applyInvM(rk, zk);
  pk = zk;
  const double rk0norm = std::sqrt(dot(rk, rk));
  double zk_rk = dot(zk, rk), rel = 0.0;
  int k = 0;
  while (true) {
    applyA(pk, Apk);
    const double alpha = zk_rk / dot(pk, Apk);
    for (int64_t i = 0; i < n; i++) { xk[i] += pk[i] * alpha; rk[i] -= Apk[i] * alpha; }
    rel = std::sqrt(dot(rk, rk)) / rk0norm;
    if (rel < desired) break;
    if (k + 1 >= maxIterations) break;
    applyInvM(rk, zk1);
    const double zk1_rk1 = dot(zk1, rk);
    const double beta = zk1_rk1 / zk_rk;
    for (int64_t i = 0; i < n; i++) pk[i] = zk1[i] + pk[i] * beta;
    zk_rk = zk1_rk1;
    k++;
  }
  x = xk;
  return {k + 1, rel};
}

// BlockJacobiPrecond::init (Preconditioner.h:71-82): in-place LLT of one d x d diagonal block (lower, row-major)
inline void jacobiFactorBlock(double* L, int d) {
  for (int a = 0; a < d; a++) {
    for (int b = 0; b < a; b++) { double t = L[a * d + b]; for (int k = 0; k < b; k++) t -= L[a * d + k] * L[b * d + k]; L[a * d + b] = t / L[b * d + b]; }
    double t = L[a * d + a];
    for (int k = 0; k < a; k++) t -= L[a * d + k] * L[a * d + k];
    if (!(t > 0)) throw std::runtime_error("oracle: Jacobi block not positive definite");
    L[a * d + a] = std::sqrt(t);
  }
}
// BlockJacobiPrecond::operator() (:84-107): x <- L^-T L^-1 x on one block
inline void jacobiSolveBlock(const double* L, int d, double* x) {
  for (int a = 0; a < d; a++) { double t = x[a]; for (int k = 0; k < a; k++) t -= L[a * d + k] * x[k]; x[a] = t / L[a * d + a]; }
  for (int a = d - 1; a >= 0; a--) { double t = x[a]; for (int k = a + 1; k < d; k++) t -= L[k * d + a] * x[k]; x[a] = t / L[a * d + a]; }
}

// PCG on a dense symmetric system (row-major n x n) with the preconditioners of Preconditioner.h over the given block
// sizes — 0 identity, 1 block Jacobi, 2 symmetric block Gauss-Seidel (D + E) D^-1 (D + E)^T, 3 lower-precision solve
// (float Cholesky of the whole system, applied in float): the reduced system of TestPCG.cpp's recipe.
inline PcgResult pcgDense(int64_t n, const double* A, const double* b, const int32_t* blockSize, int64_t nBlocks, int precond,
                          double desired, int maxIterations, double* x) {
  std::vector<int64_t> start((size_t)nBlocks + 1, 0), off((size_t)nBlocks + 1, 0);
  for (int64_t c = 0; c < nBlocks; c++) { start[c + 1] = start[c] + blockSize[c]; off[c + 1] = off[c] + (int64_t)blockSize[c] * blockSize[c]; }
  if (start[nBlocks] != n) throw std::runtime_error("oracle: block sizes do not add up to the system size");
  std::vector<double> jac((size_t)off[nBlocks], 0.0);
  if (precond == 1 || precond == 2)
    for (int64_t c = 0; c < nBlocks; c++) {
      const int d = blockSize[c];
      double* L = &jac[off[c]];
      for (int a = 0; a < d; a++) for (int k = 0; k <= a; k++) L[a * d + k] = A[(start[c] + a) * n + start[c] + k];
      jacobiFactorBlock(L, d);
    }
  std::vector<float> Lf;
  if (precond == 3) {   // LowerPrecSolvePrecond::init: raise the diagonal until the float factor is finite
    float epsilon = 0.0f;
    while (true) {
      Lf.assign(A, A + n * n);
      if (epsilon > 0) { for (int64_t i = 0; i < n; i++) { Lf[i * n + i] *= 1.0f + epsilon; Lf[i * n + i] += epsilon; } epsilon *= 3.0f; }
      else epsilon = 1e-8f;
      bool finite = true;
      for (int64_t i = 0; i < n && finite; i++) {
        for (int64_t j = 0; j < i; j++) { float t = Lf[i * n + j]; for (int64_t k = 0; k < j; k++) t -= Lf[i * n + k] * Lf[j * n + k]; Lf[i * n + j] = t / Lf[j * n + j]; }
        float d = Lf[i * n + i];
        for (int64_t k = 0; k < i; k++) d -= Lf[i * n + k] * Lf[i * n + k];
        Lf[i * n + i] = std::sqrt(d);
        finite = std::isfinite(Lf[i * n + i]) && Lf[i * n + i] > 0.0f;
      }
      if (finite) break;
      if (epsilon > 1e3f) throw std::runtime_error("oracle: lower-precision factor did not become finite");
    }
  }
  std::vector<double> rhs(b, b + n), sol;
  const PcgResult r = pcgRun(
      [&](const std::vector<double>& v, std::vector<double>& u) {
        u.assign((size_t)n, 0.0);
        for (int64_t i = 0; i < n; i++) { double s = 0; const double* Ai = A + i * n; for (int64_t j = 0; j < n; j++) s += Ai[j] * v[j]; u[i] = s; }
      },
      [&](const std::vector<double>& rr, std::vector<double>& z) {
        z = rr;
        if (precond == 1) for (int64_t c = 0; c < nBlocks; c++) jacobiSolveBlock(&jac[off[c]], blockSize[c], &z[start[c]]);
        if (precond == 2) {
          for (int64_t c = 0; c < nBlocks; c++) {                       // (D + E) w = r
            for (int64_t i = start[c]; i < start[c + 1]; i++) { double t = z[i]; for (int64_t j = 0; j < start[c]; j++) t -= A[i * n + j] * z[j]; z[i] = t; }
            jacobiSolveBlock(&jac[off[c]], blockSize[c], &z[start[c]]);
          }
          for (int64_t c = 0; c < nBlocks; c++) {                       // u = D w
            double u[64];
            for (int64_t i = start[c]; i < start[c + 1]; i++) { double t = 0; for (int64_t j = start[c]; j < start[c + 1]; j++) t += A[i * n + j] * z[j]; u[i - start[c]] = t; }
            for (int64_t i = start[c]; i < start[c + 1]; i++) z[i] = u[i - start[c]];
          }
          for (int64_t c = nBlocks - 1; c >= 0; c--) {                  // (D + E)^T z = u
            jacobiSolveBlock(&jac[off[c]], blockSize[c], &z[start[c]]);
            for (int64_t i = start[c]; i < start[c + 1]; i++) for (int64_t j = 0; j < start[c]; j++) z[j] -= A[i * n + j] * z[i];
          }
        }
        if (precond == 3) {
          std::vector<float> y(rr.begin(), rr.end());
          for (int64_t i = 0; i < n; i++) { float t = y[i]; for (int64_t k = 0; k < i; k++) t -= Lf[i * n + k] * y[k]; y[i] = t / Lf[i * n + i]; }
          for (int64_t i = n - 1; i >= 0; i--) { float t = y[i]; for (int64_t k = i + 1; k < n; k++) t -= Lf[k * n + i] * y[k]; y[i] = t / Lf[i * n + i]; }
          for (int64_t i = 0; i < n; i++) z[i] = (double)y[i];
        }
      },
      rhs, desired, maxIterations, sol);
  for (int64_t i = 0; i < n; i++) x[i] = sol[i];
  return r;
}

}  // namespace orc
