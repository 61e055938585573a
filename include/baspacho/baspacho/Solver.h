// baspacho/baspacho/Solver.h — facade of BaSpaCho::Solver for the calls of the reference's DIRECT path:
//   Solver::order(), dataSize()                       lib/small_thing/Optimizer.cpp:339-340,777,783-784
//   Solver::accessor()                                Factor.h:612,633,643; Variable.h:360; Optimizer.cpp:142
//   Solver::factor(data)                              Optimizer.cpp:223  (landmark Schur elimination + supernodal Cholesky)
//   Solver::solve(data, vec, stride, nRHS)            Optimizer.cpp:228
// over a device problem of this library (the C ABI of viba_b200.h): `data` and `vec` are HOST arrays in this library's
// layout (dataSize() / order() doubles, permuted order), moved to the device around the call; the factor stays on the
// device between factor() and solve(), and factor() also writes the factored panels back into `data` (in place, as
// BaSpaCho's contract says; the bytes are this library's panels, not BaSpaCho's).
//
// Not offered (not on ark_vi_ba's direct path): createSolver from a bare block structure, the partial
// factor / solve family (factorUpTo, solveLUpTo, ..., addMvFrom), skel().  They are declared so that code naming them
// still compiles, and throw.  The reference's own BaSpaCho source is absent from its tree (deps/baspacho is an
// un-vendored submodule pinned at bd833260082023d5581b4996360a43a4de058d49), so this surface is reconstructed from the
// reference's call sites only.
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../viba_b200.h"
#include "Accessor.h"
#include "DebugMacros.h"

namespace BaSpaCho {

enum AddFillPolicy { AddFillComplete, AddFillForGivenElims, AddFillForAutoElims };
enum BackendType { BackendRef, BackendFast, BackendCuda };

struct Settings {
  bool findSparseEliminationRanges = true;
  int numThreads = 16;
  BackendType backend = BackendCuda;   // informational: the numeric phase always runs on the sm_100a device
  AddFillPolicy addFillPolicy = AddFillComplete;
};

struct SparseStructure {   // lower-triangular block CSR including the diagonal (Optimizer.cpp:174-197)
  std::vector<int64_t> ptrs, inds;
  SparseStructure() = default;
  SparseStructure(std::vector<int64_t> p, std::vector<int64_t> i) : ptrs(std::move(p)), inds(std::move(i)) {}
};

class Solver {
 public:
  // wraps a problem created through the C ABI (the small_thing mirror hands its backend: Optimizer::backendProblem())
  explicit Solver(vb_problem* p) : p_(p), acc_(p) { BASPACHO_CHECK_NOTNULL(p); }

  int64_t order() const { return vb_order(p_); }
  int64_t dataSize() const { return vb_data_size(p_); }
  const PermutedCoalescedAccessor& accessor() const { return acc_; }

  // in-place factorisation of the flat array (Optimizer.cpp:223)
  void factor(double* data) const {
    check(vb_upload_hess(p_, data, dataSize()));
    check(vb_factor(p_));
    check(vb_download_hess(p_, data, dataSize()));
  }
  // in-place solve with the factor computed by the last factor() (Optimizer.cpp:228): vec holds nRHS columns of
  // `stride` doubles; `data` must be the array factor() was given (it is not re-read: the factor lives on the device)
  void solve(const double* data, double* vec, int64_t stride, int nRHS) const {
    (void)data;
    BASPACHO_CHECK_GE(stride, order());
    for (int r = 0; r < nRHS; r++) {
      double* v = vec + (int64_t)r * stride;
      check(vb_upload_grad(p_, v, order()));
      check(vb_step_from_grad(p_));
      check(vb_solve(p_, 0));
      check(vb_download_step(p_, v, order()));
    }
  }

  // ---- the partial family and the skeleton: named by lib/small_thing (PCG, marginals, covariances), not reached by
  // ark_vi_ba's direct path; declared for source compatibility
  template <typename T> void factorUpTo(T*, int64_t) const { unsupported("factorUpTo"); }
  template <typename T> void factorFrom(T*, int64_t) const { unsupported("factorFrom"); }
  template <typename T> void solveLUpTo(const T*, int64_t, T*, int64_t, int) const { unsupported("solveLUpTo"); }
  template <typename T> void solveLtUpTo(const T*, int64_t, T*, int64_t, int) const { unsupported("solveLtUpTo"); }
  template <typename T> void solveLFrom(const T*, int64_t, T*, int64_t, int) const { unsupported("solveLFrom"); }
  template <typename T> void solveLtFrom(const T*, int64_t, T*, int64_t, int) const { unsupported("solveLtFrom"); }
  template <typename T> void pseudoFactorFrom(T*, int64_t) const { unsupported("pseudoFactorFrom"); }
  template <typename T> void addMvFrom(const T*, int64_t, const T*, int64_t, T*, int64_t, int, T = T(1)) const { unsupported("addMvFrom"); }

 private:
  static void check(int rc) {
    if (rc != VB_OK) throw std::runtime_error(std::string("BaSpaCho facade: ") + vb_last_error());
  }
  [[noreturn]] static void unsupported(const char* what) {
    throw std::runtime_error(std::string("BaSpaCho facade: Solver::") + what + " is not on the device path (SURVEY.md §8 f3/f4)");
  }
  vb_problem* p_;
  PermutedCoalescedAccessor acc_;
};

using SolverPtr = std::unique_ptr<Solver>;

// createSolver from a bare block structure (Optimizer.cpp:200-206) needs a symbolic phase without factor tables, which
// this library derives from the factor stores instead (vb_problem_create): not offered.
inline SolverPtr createSolver(const Settings&, const std::vector<int64_t>&, const SparseStructure&, const std::vector<int64_t>& = {},
                              const std::vector<int64_t>& = {}) {
  throw std::runtime_error("BaSpaCho facade: createSolver(structure) is not offered; wrap a vb_problem (small_thing::Optimizer::backendProblem())");
}

}  // namespace BaSpaCho
