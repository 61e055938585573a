// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Stand-in for BaSpaCho::Solver::factor + solve as consumed at lib/small_thing/Optimizer.cpp:223,228
// (BaSpaCho facebookresearch/baspacho@bd833260 is NOT in /root/reference: SURVEY.md F1).  The
// published algorithm is restated in its simplest exact form: eliminate the landmark range
// [0,P) block by block (3x3 Cholesky, Schur complement onto the camera-side blocks), then a
// direct Cholesky of the reduced camera system, here an envelope (skyline) Cholesky under a
// reverse Cuthill-McKee block ordering.  Ordering / supernodes / data layout: PARITY UNPINNED
// (no reference source or golden file fixes them); the solution vector is ordering independent
// up to rounding, which is what tests compare (rel 1e-9).
#pragma once
#include "pcg.hpp"
#include <queue>
#include <numeric>
#include "engine.hpp"

namespace orc {

struct SchurSkylineSolver {
  const Engine* E = nullptr;
  int64_t nP = 0, nB = 0, nC = 0;  // landmark blocks, all blocks, camera-side blocks
  int64_t nr = 0;                  // reduced scalar order
  std::vector<std::vector<int64_t>> colBlocks; // per landmark: CSR entry ids k of blocks (j,p)
  std::vector<int64_t> cpos, cstart;           // camera block (reg index - nP) -> position / scalar start in reduced order
  std::vector<int64_t> first, rowStart;        // skyline per scalar row
  std::vector<double> S, gr;                   // skyline data, reduced rhs
  std::vector<double> HppInv;                  // 9 per landmark (inverse of damped H_pp)
  double flopsSchur = 0, flopsChol = 0;

  void analyze(const Engine& eng) {
    E = &eng;
    nP = eng.P.npoints; nB = eng.P.nblocks; nC = nB - nP;
    colBlocks.assign(nP, {});
    std::vector<std::vector<int64_t>> adj(nC);
    for (int64_t i = nP; i < nB; i++)
      for (int64_t k = eng.rowPtr[i]; k < eng.rowPtr[i + 1] - 1; k++) {
        const int64_t j = eng.colInd[k];
        if (j < nP) colBlocks[j].push_back(k);
        else { adj[i - nP].push_back(j - nP); adj[j - nP].push_back(i - nP); }
      }
    for (int64_t p = 0; p < nP; p++)
      if (eng.rowPtr[p + 1] - eng.rowPtr[p] != 1) throw std::runtime_error("oracle: landmark-landmark block");
    // row of each CSR entry
    std::vector<int64_t> rowOf(eng.colInd.size());
    for (int64_t i = 0; i < nB; i++) for (int64_t k = eng.rowPtr[i]; k < eng.rowPtr[i + 1]; k++) rowOf[k] = i;
    rowOfEntry = rowOf;
    for (int64_t p = 0; p < nP; p++) {
      const auto& cb = colBlocks[p];
      for (size_t a = 0; a < cb.size(); a++)
        for (size_t b = 0; b < a; b++) {
          const int64_t ia = rowOf[cb[a]] - nP, ib = rowOf[cb[b]] - nP;
          adj[ia].push_back(ib); adj[ib].push_back(ia);
        }
    }
    for (auto& a : adj) { std::sort(a.begin(), a.end()); a.erase(std::unique(a.begin(), a.end()), a.end()); }
    // reverse Cuthill-McKee over all components, seeds = min degree unvisited
    orderC.clear(); orderC.reserve(nC);
    std::vector<char> vis(nC, 0);
    std::vector<int64_t> byDeg(nC);
    std::iota(byDeg.begin(), byDeg.end(), 0);
    std::stable_sort(byDeg.begin(), byDeg.end(), [&](int64_t a, int64_t b) { return adj[a].size() < adj[b].size(); });
    for (int64_t seed : byDeg) {
      if (vis[seed]) continue;
      std::queue<int64_t> q; q.push(seed); vis[seed] = 1;
      while (!q.empty()) {
        const int64_t u = q.front(); q.pop(); orderC.push_back(u);
        std::vector<int64_t> nb;
        for (int64_t w : adj[u]) if (!vis[w]) { vis[w] = 1; nb.push_back(w); }
        std::stable_sort(nb.begin(), nb.end(), [&](int64_t a, int64_t b) { return adj[a].size() < adj[b].size(); });
        for (int64_t w : nb) q.push(w);
      }
    }
    std::reverse(orderC.begin(), orderC.end());
    cpos.assign(nC, 0); cstart.assign(nC, 0);
    for (int64_t i = 0; i < nC; i++) cpos[orderC[i]] = i;
    nr = 0;
    for (int64_t i = 0; i < nC; i++) { cstart[orderC[i]] = nr; nr += eng.P.blockSize[nP + orderC[i]]; }
    // envelope
    first.assign(nr, 0);
    for (int64_t c = 0; c < nC; c++) {
      int64_t f = cstart[c];
      for (int64_t w : adj[c]) f = std::min(f, cstart[w]);
      for (int a = 0; a < eng.P.blockSize[nP + c]; a++) first[cstart[c] + a] = f;
    }
    rowStart.assign(nr + 1, 0);
    for (int64_t i = 0; i < nr; i++) rowStart[i + 1] = rowStart[i] + (i - first[i] + 1);
    S.assign(rowStart[nr], 0.0);
    gr.assign(nr, 0.0);
    HppInv.assign(9 * nP, 0.0);
    flopsChol = 0;
    for (int64_t i = 0; i < nr; i++) { const double w = double(i - first[i]); flopsChol += w * w; }
  }
  std::vector<int64_t> rowOfEntry;
  std::vector<int64_t> orderC;   // camera blocks in the order of the reduced system (reverse Cuthill-McKee)

  inline double& at(int64_t i, int64_t j) { return S[rowStart[i] + (j - first[i])]; } // i >= j >= first[i]

  // add block (reg indices bi >= bj, camera side) d_i x d_j row-major with sign
  void addBlock(int64_t bi, int64_t bj, const double* B, double sign) {
    const int di = E->P.blockSize[bi], dj = E->P.blockSize[bj];
    const int64_t si = cstart[bi - nP], sj = cstart[bj - nP];
    const bool par = E->threads > 1; // CPU-baseline mode: landmarks are eliminated by several threads
    auto add = [par](double& dst, double v) {
      if (par) {
#pragma omp atomic
        dst += v;
      } else dst += v;
    };
    if (bi == bj) {
      for (int a = 0; a < di; a++) for (int b = 0; b <= a; b++) add(at(si + a, sj + b), sign * B[a * dj + b]);
    } else if (si > sj) {
      for (int a = 0; a < di; a++) for (int b = 0; b < dj; b++) add(at(si + a, sj + b), sign * B[a * dj + b]);
    } else {
      for (int a = 0; a < di; a++) for (int b = 0; b < dj; b++) add(at(sj + b, si + a), sign * B[a * dj + b]);
    }
  }
  // ---- CPU-baseline mode only: LAPACK band Cholesky (dpbtrf/dpbtrs) of the reduced system on a
  // multi-threaded OpenBLAS loaded at run time (the reference factors with BaSpaCho's BLAS backend on
  // OpenBLAS, README.md:109-118).  Under the RCM ordering the envelope is a band, so the blocked
  // BLAS-3 band routine is the closest stock stand-in for a supernodal factorisation.
  typedef void (*pbtrf_fn)(const char*, const int*, const int*, double*, const int*, int*);
  typedef void (*pbtrs_fn)(const char*, const int*, const int*, const int*, const double*, const int*, double*, const int*, int*);
  pbtrf_fn pbtrf = nullptr; pbtrs_fn pbtrs = nullptr;
  std::vector<double> AB; int kd = 0; bool bandFactored = false;

  static void inv3sym(const double* A, double* I) { // symmetric 3x3 (lower used), via Cholesky
    const double l00 = std::sqrt(A[0]), l10 = A[3] / l00, l20 = A[6] / l00;
    const double l11 = std::sqrt(A[4] - l10 * l10), l21 = (A[7] - l20 * l10) / l11;
    const double l22 = std::sqrt(A[8] - l20 * l20 - l21 * l21);
    // inverse of L
    const double i00 = 1 / l00, i11 = 1 / l11, i22 = 1 / l22;
    const double i10 = -l10 * i00 * i11;
    const double i21 = -l21 * i11 * i22;
    const double i20 = -(l20 * i00 + l21 * i10) * i22;
    // A^-1 = L^-T L^-1
    I[0] = i00 * i00 + i10 * i10 + i20 * i20; I[1] = i10 * i11 + i20 * i21; I[2] = i20 * i22;
    I[3] = I[1]; I[4] = i11 * i11 + i21 * i21; I[5] = i21 * i22;
    I[6] = I[2]; I[7] = I[5]; I[8] = i22 * i22;
  }

  // factor(hess) + solve(hess, g) : x = H^-1 g, all in registration order
  void factorAndSolve(const double* H, const double* g, double* x) { factor(H); solveOnly(H, g, x); }

  // Solver::factor (Optimizer.cpp:223): H is read-only here; the factor lives in S / HppInv
  void factor(const double* H) {
    const Engine& e = *E;
    std::fill(S.begin(), S.end(), 0.0);
    // camera-side blocks
    for (int64_t i = nP; i < nB; i++)
      for (int64_t k = e.rowPtr[i]; k < e.rowPtr[i + 1]; k++) {
        const int64_t j = e.colInd[k];
        if (j >= nP) addBlock(i, j, H + e.blkOff[k], 1.0);
      }
    // landmark elimination
    flopsSchur = 0;
    double fl = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : fl) num_threads(e.threads) if (e.threads > 1)
    for (int64_t p = 0; p < nP; p++) {
      std::vector<double> Y;
      const double* Hpp = H + e.blkOff[e.rowPtr[p + 1] - 1];
      double* Hi = &HppInv[9 * p];
      inv3sym(Hpp, Hi);
      const auto& cb = colBlocks[p];
      int64_t m = 0;
      for (int64_t k : cb) m += e.P.blockSize[rowOfEntry[k]];
      fl += 3.0 * m * m + 9.0 * m + 9.0;
      // Y_a = W_a Hpp^-1 (d_a x 3), W_a = block (a,p) (d_a x 3)
      Y.assign(3 * m, 0.0);
      int64_t o = 0;
      for (int64_t k : cb) {
        const int da = e.P.blockSize[rowOfEntry[k]];
        matmul(H + e.blkOff[k], Hi, &Y[3 * o], da, 3, 3);
        o += da;
      }
      int64_t oa = 0;
      double blk[23 * 23];
      for (size_t a = 0; a < cb.size(); a++) {
        const int64_t ba = rowOfEntry[cb[a]];
        const int da = e.P.blockSize[ba];
        const double* Wa = H + e.blkOff[cb[a]];
        (void)Wa;
        int64_t ob = 0;
        for (size_t b = 0; b <= a; b++) {
          const int64_t bb = rowOfEntry[cb[b]];
          const int db = e.P.blockSize[bb];
          const double* Wb = H + e.blkOff[cb[b]];
          // Y_a W_b^T : da x db
          for (int s = 0; s < da; s++) for (int t = 0; t < db; t++) blk[s * db + t] = dot3(&Y[3 * (oa + s)], Wb + 3 * t);
          addBlock(ba, bb, blk, -1.0); // ba > bb since CSR rows ascend
          ob += db;
        }
        oa += da;
      }
    }
    flopsSchur = fl;
    bandFactored = false;
    factoredDirect = false;
    if (schurOnly) { initPrecond(); return; }   // iterative path (Optimizer.cpp:291-292: factorUpTo the elimination range only)
    factoredDirect = true;
    if (pbtrf && e.threads > 1) { // CPU-baseline mode: band storage (lower), AB[(i-j) + j*ldab] = A(i,j)
      kd = 0;
      for (int64_t i = 0; i < nr; i++) kd = std::max<int>(kd, (int)(i - first[i]));
      const int64_t ldab = kd + 1;
      AB.assign((size_t)ldab * nr, 0.0);
#pragma omp parallel for schedule(static) num_threads(e.threads)
      for (int64_t i = 0; i < nr; i++)
        for (int64_t j = first[i]; j <= i; j++) AB[(size_t)(i - j) + (size_t)j * ldab] = S[rowStart[i] + (j - first[i])];
      const int n = (int)nr, ld = (int)ldab;
      int info = 0;
      pbtrf("L", &n, &kd, AB.data(), &ld, &info);
      if (info != 0) throw std::runtime_error("oracle: dpbtrf failed (reduced system not positive definite)");
      bandFactored = true;
      return;
    }
    // envelope Cholesky (row by row)
    for (int64_t i = 0; i < nr; i++) {
      double* Li = &S[rowStart[i]];
      const int64_t fi = first[i];
      for (int64_t j = fi; j < i; j++) {
        const double* Lj = &S[rowStart[j]];
        const int64_t fj = first[j];
        const int64_t k0 = std::max(fi, fj);
        double s = Li[j - fi];
        const double* a = Li + (k0 - fi);
        const double* b = Lj + (k0 - fj);
        const int64_t n = j - k0;
        for (int64_t k = 0; k < n; k++) s -= a[k] * b[k];
        Li[j - fi] = s / Lj[j - fj];
      }
      double d = Li[i - fi];
      for (int64_t k = 0; k < i - fi; k++) d -= Li[k] * Li[k];
      if (!(d > 0)) throw std::runtime_error("oracle: reduced system not positive definite");
      Li[i - fi] = std::sqrt(d);
    }
  }

  // ---- iterative path: Optimizer.cpp:232-331 (PCG on the reduced camera system), PCG.cpp:15-97, and the two
  // Preconditioners (Preconditioner.h:29-245).  IdentityPrecond and BlockJacobiPrecond (the param-sized diagonal blocks,
  // Cholesky-factored) are defined in the reference tree.  BlockGaussSeidelPrecond is Solver::pseudoFactorFrom +
  // solveLFrom + solveLtFrom of BaSpaCho (source absent): restated here from its call site, its name and its comment
  // ("param-sized blocks are inverted") as the symmetric block Gauss-Seidel operator M = (D + E) D^-1 (D + E)^T over the
  // param blocks (D block diagonal, E strict block lower part) — the operator a Cholesky sweep without Schur updates
  // yields — and checked against the bound the reference's own test puts on it (TestPCG.cpp:139-141; parity unpinned
  // beyond that).  LowerPrecSolvePrecond (:160-245) is fully defined by its call site: the Cholesky factor of the float
  // copy of the system, diagonal raised by d -> d (1 + eps) + eps, eps = 1e-8, 3e-8, ... while the factor is not finite,
  // applied in float.
  bool schurOnly = false, factoredDirect = false;
  int pcgPrecond = 0;            // 0 identity, 1 block Jacobi, 2 block Gauss-Seidel, 3 lower-precision solve
  int pcgMaxIterations = 40; double pcgDesiredResidual = 1e-10;
  int pcgLastIterations = 0; double pcgLastResidual = 0;
  // Optimizer::Settings::solverType (Optimizer.h:30-36,43): 0 direct, 1 PCG identity, 2 Jacobi, 3 Gauss-Seidel, 4 lower-prec
  void setLinearSolver(int solverType, int maxIterations, double desiredResidual) {
    if (solverType < 0 || solverType > 4) throw std::runtime_error("Unknown preconditioner " + std::to_string(solverType));   // Optimizer.cpp:243-245
    schurOnly = solverType != 0; pcgPrecond = solverType > 0 ? solverType - 1 : 0;
    pcgMaxIterations = maxIterations; pcgDesiredResidual = desiredResidual;
  }
  std::vector<float> Sf;         // lower-precision solve: float envelope Cholesky factor of the reduced system
  float lowPrecEpsilon = 0;      // the diagonal raise its last init ended with (0: none was needed)
  std::vector<double> jac;       // per camera block: Cholesky factor of its diagonal block (d x d, lower), at jacOff
  std::vector<int64_t> jacOff;

  void reducedMatVec(const std::vector<double>& v, std::vector<double>& u) const {   // u = S v, S symmetric in skyline
    u.assign(nr, 0.0);
    for (int64_t i = 0; i < nr; i++) {
      const double* Si = &S[rowStart[i]];
      const int64_t fi = first[i];
      double s = 0;
      for (int64_t j = fi; j < i; j++) { s += Si[j - fi] * v[j]; u[j] += Si[j - fi] * v[i]; }
      u[i] += s + Si[i - fi] * v[i];
    }
  }
  void initPrecond() {
    const Engine& e = *E;
    jacOff.assign(nC + 1, 0);
    for (int64_t c = 0; c < nC; c++) jacOff[c + 1] = jacOff[c] + (int64_t)e.P.blockSize[nP + c] * e.P.blockSize[nP + c];
    jac.assign(jacOff[nC], 0.0);
    if (pcgPrecond == 3) { initLowPrec(); return; }
    if (pcgPrecond != 1 && pcgPrecond != 2) return;
    for (int64_t c = 0; c < nC; c++) {
      const int d = e.P.blockSize[nP + c];
      double* L = &jac[jacOff[c]];
      for (int a = 0; a < d; a++) for (int b = 0; b <= a; b++) L[a * d + b] = S[rowStart[cstart[c] + a] + (cstart[c] + b - first[cstart[c] + a])];
      jacobiFactorBlock(L, d);
    }
  }
  // LowerPrecSolvePrecond::init (Preconditioner.h:178-218)
  void initLowPrec() {
    float epsilon = 0.0f;
    lowPrecEpsilon = 0.0f;
    while (true) {
      Sf.assign(S.begin(), S.end());
      if (epsilon > 0) {
        for (int64_t i = 0; i < nr; i++) { float& d = Sf[rowStart[i] + (i - first[i])]; d *= 1.0f + epsilon; d += epsilon; }
        lowPrecEpsilon = epsilon;
        epsilon *= 3.0f;
      } else {
        epsilon = 1e-8f;
      }
      bool finite = true;
      for (int64_t i = 0; i < nr && finite; i++) {   // the envelope Cholesky of factor(), in float
        float* Li = &Sf[rowStart[i]];
        const int64_t fi = first[i];
        for (int64_t j = fi; j < i; j++) {
          const float* Lj = &Sf[rowStart[j]];
          const int64_t fj = first[j], k0 = std::max(fi, fj);
          float t = Li[j - fi];
          for (int64_t k = k0; k < j; k++) t -= Li[k - fi] * Lj[k - fj];
          Li[j - fi] = t / Lj[j - fj];
        }
        float d = Li[i - fi];
        for (int64_t k = 0; k < i - fi; k++) d -= Li[k] * Li[k];
        Li[i - fi] = std::sqrt(d);
        finite = std::isfinite(Li[i - fi]) && Li[i - fi] > 0.0f;
      }
      if (finite) return;
      if (epsilon > 1e3f) throw std::runtime_error("oracle: lower-precision factor did not become finite");
    }
  }
  void applyPrecond(const std::vector<double>& r, std::vector<double>& z) const {
    z = r;
    const Engine& e = *E;
    if (pcgPrecond == 2) {   // (D + E) w = r ; u = D w ; (D + E)^T z = u
      // blocks in the order of the reduced system (orderC), not in registration order
      for (int64_t oc = 0; oc < nC; oc++) {
        const int64_t c = orderC[oc];
        const int d = e.P.blockSize[nP + c];
        for (int a = 0; a < d; a++) {
          const int64_t i = cstart[c] + a;
          const double* Si = &S[rowStart[i]];
          double t = z[i];
          for (int64_t j = first[i]; j < cstart[c]; j++) t -= Si[j - first[i]] * z[j];
          z[i] = t;
        }
        jacobiSolveBlock(&jac[jacOff[c]], d, &z[cstart[c]]);
      }
      for (int64_t c = 0; c < nC; c++) {   // u = D w
        const int d = e.P.blockSize[nP + c];
        double w[32], u[32];
        for (int a = 0; a < d; a++) w[a] = z[cstart[c] + a];
        for (int a = 0; a < d; a++) {
          const int64_t i = cstart[c] + a;
          double t = 0;
          for (int b = 0; b < d; b++) {
            const int64_t j = cstart[c] + b;
            t += (b <= a ? S[rowStart[i] + (j - first[i])] : S[rowStart[j] + (i - first[j])]) * w[b];
          }
          u[a] = t;
        }
        for (int a = 0; a < d; a++) z[cstart[c] + a] = u[a];
      }
      for (int64_t oc = nC - 1; oc >= 0; oc--) {
        const int64_t c = orderC[oc];
        const int d = e.P.blockSize[nP + c];
        jacobiSolveBlock(&jac[jacOff[c]], d, &z[cstart[c]]);
        for (int a = 0; a < d; a++) {
          const int64_t i = cstart[c] + a;
          const double* Si = &S[rowStart[i]];
          const double zi = z[i];
          for (int64_t j = first[i]; j < cstart[c]; j++) z[j] -= Si[j - first[i]] * zi;
        }
      }
      return;
    }
    if (pcgPrecond == 3) {   // LowerPrecSolvePrecond::operator() (:220-239): cast, two float triangular solves, cast back
      std::vector<float> y(r.begin(), r.end());
      for (int64_t i = 0; i < nr; i++) {
        const float* Li = &Sf[rowStart[i]];
        float t = y[i];
        for (int64_t k = first[i]; k < i; k++) t -= Li[k - first[i]] * y[k];
        y[i] = t / Li[i - first[i]];
      }
      for (int64_t i = nr - 1; i >= 0; i--) {
        const float* Li = &Sf[rowStart[i]];
        y[i] /= Li[i - first[i]];
        const float yi = y[i];
        for (int64_t k = first[i]; k < i; k++) y[k] -= Li[k - first[i]] * yi;
      }
      for (int64_t i = 0; i < nr; i++) z[i] = (double)y[i];
      return;
    }
    if (pcgPrecond != 1) return;
    for (int64_t c = 0; c < nC; c++) jacobiSolveBlock(&jac[jacOff[c]], e.P.blockSize[nP + c], &z[cstart[c]]);
  }
  // PCG::solve (PCG.cpp:15-97): x0 = 0, stops on |r_k| / |r_0| < desired or at maxIterations
  void pcgSolve(const std::vector<double>& b, std::vector<double>& x) {
    const PcgResult r = pcgRun([&](const std::vector<double>& v, std::vector<double>& u) { reducedMatVec(v, u); },
                               [&](const std::vector<double>& rr, std::vector<double>& z) { applyPrecond(rr, z); },
                               b, pcgDesiredResidual, pcgMaxIterations, x);
    pcgLastIterations = r.numIterations; pcgLastResidual = r.relativeResidual;
  }

  // Solver::solve (Optimizer.cpp:228) with the factor computed by factor(H)
  void solveOnly(const double* H, const double* g, double* x) {
    const Engine& e = *E;
    for (int64_t c = 0; c < nC; c++)
      for (int a = 0; a < e.P.blockSize[nP + c]; a++) gr[cstart[c] + a] = g[e.paramStart[nP + c] + a];
    for (int64_t p = 0; p < nP; p++) {
      double hg[3];
      mat3vec(&HppInv[9 * p], g + e.paramStart[p], hg);
      for (int64_t k : colBlocks[p]) {
        const int64_t ba = rowOfEntry[k];
        const int da = e.P.blockSize[ba];
        const double* Wa = H + e.blkOff[k];
        for (int t = 0; t < da; t++) gr[cstart[ba - nP] + t] -= dot3(Wa + 3 * t, hg);
      }
    }
    // solve L y = gr ; L^T z = y
    std::vector<double> y(gr);
    if (!factoredDirect) {   // iterative path: PCG on the reduced system (Optimizer.cpp:308-310)
      std::vector<double> sol;
      pcgSolve(gr, sol);
      y = sol;
    } else if (bandFactored) {
      const int n = (int)nr, ld = kd + 1, one = 1;
      int info = 0;
      pbtrs("L", &n, &kd, &one, AB.data(), &ld, y.data(), &n, &info);
      if (info != 0) throw std::runtime_error("oracle: dpbtrs failed");
    } else {
    for (int64_t i = 0; i < nr; i++) {
      const double* Li = &S[rowStart[i]];
      double s = y[i];
      for (int64_t k = first[i]; k < i; k++) s -= Li[k - first[i]] * y[k];
      y[i] = s / Li[i - first[i]];
    }
    for (int64_t i = nr - 1; i >= 0; i--) {
      const double* Li = &S[rowStart[i]];
      y[i] /= Li[i - first[i]];
      const double yi = y[i];
      for (int64_t k = first[i]; k < i; k++) y[k] -= Li[k - first[i]] * yi;
    }
    }
    for (int64_t c = 0; c < nC; c++)
      for (int a = 0; a < e.P.blockSize[nP + c]; a++) x[e.paramStart[nP + c] + a] = y[cstart[c] + a];
    // back-substitute landmarks: x_p = Hpp^-1 (g_p - sum_a W_a^T x_a)
    for (int64_t p = 0; p < nP; p++) {
      double rhs[3] = {g[e.paramStart[p]], g[e.paramStart[p] + 1], g[e.paramStart[p] + 2]};
      for (int64_t k : colBlocks[p]) {
        const int64_t ba = rowOfEntry[k];
        const int da = e.P.blockSize[ba];
        const double* Wa = H + e.blkOff[k];
        const double* xa = x + e.paramStart[ba];
        for (int t = 0; t < da; t++) for (int c = 0; c < 3; c++) rhs[c] -= Wa[3 * t + c] * xa[t];
      }
      mat3vec(&HppInv[9 * p], rhs, x + e.paramStart[p]);
    }
  }
};

} // namespace orc
