// Host-side symbolic phase (one-off per problem): replaces Optimizer::initSolver
// (lib/small_thing/Optimizer.cpp:166-207), FactorStore::registerBlocks (Factor.h:761-785) and the
// symbolic half of BaSpaCho::createSolver (source absent from the reference tree: SURVEY.md F1,
// §2.3 — ordering / supernode partition / layout are this library's own, "parity unpinned").
//
// Layout of the flat Hessian array ("hess"), all FP64, every panel row-major:
//   landmark p (registration index p < nP, eliminated first, elimRanges=[0,nP)):
//     panel of (3 + m_p) rows x 3 cols at lmOff[p]: H_pp, then one d_a x 3 block per camera-side
//     neighbour a, ascending elimination position.
//   camera-side lump L (supernode = consecutive elimination positions):
//     panel of (w_L + r_L) rows x w_L cols at lumpOff[L]: the w_L x w_L diagonal part (lower
//     triangle significant) then the below-diagonal row blocks, ascending position.
// grad / step vectors are in elimination (permuted) order, landmarks first.
#pragma once
#include <cstdint>
#include <vector>
#include <string>
#include "../../include/viba_b200.h"

namespace vb {

struct BlockRef { int64_t off; int32_t stride; int32_t transposed; };

struct Symbolic {
  int64_t nB = 0, nP = 0, nC = 0, order = 0, dataSize = 0, lmDataSize = 0;
  std::vector<int32_t> bsize;     // per registration index
  std::vector<int64_t> pos;       // registration index -> elimination position
  std::vector<int64_t> posToReg;
  std::vector<int64_t> vecStart;  // [nB+1] scalar start per elimination position
  // landmark panels
  std::vector<int64_t> lmOff;     // [nP+1]
  std::vector<int64_t> lmRowPtr;  // [nP+1]
  std::vector<int32_t> lmRowCam;  // camera position (pos - nP) per row block
  std::vector<int32_t> lmRowOff;  // scalar row in the panel (>= 3)
  // camera lumps
  int64_t nL = 0;
  std::vector<int32_t> camLump;     // [nC] lump of each camera position
  std::vector<int32_t> camStart;    // [nC+1] scalar start in the reduced system
  std::vector<int32_t> lumpFirst;   // [nL+1] first camera position
  std::vector<int32_t> lumpWidth, lumpRows; // scalar cols, total rows (width + below)
  std::vector<int64_t> lumpOff;     // [nL]
  std::vector<int64_t> lumpRowPtr;  // [nL+1]
  std::vector<int32_t> lumpRowCam, lumpRowOff; // below-diagonal row blocks
  std::vector<int32_t> lumpParent, lumpLevel;
  std::vector<std::vector<int32_t>> levels;    // lumps per level
  // multi-GPU plan of the supernodal factorisation (empty / dataSize on one GPU): owner rank of every
  // lump, -1 = shared top of the elimination tree (factored by every rank); the panels of the shared
  // lumps are laid out last, from sharedOff
  std::vector<int32_t> lumpOwner;
  int64_t sharedOff = 0;
  double schurFlops = 0, cholFlops = 0;
  int32_t maxLumpWidth = 0, maxLumpRows = 0;

  // accessor().block(data, i, j) (lib/small_thing/Factor.h:633,643): registration indices, any order
  bool block(int64_t ri, int64_t rj, BlockRef* out) const;
  // row (scalar) of camera position c inside lump L's panel, or -1
  int32_t rowInLump(int32_t L, int32_t c) const;
};

struct HostGraph { // owning copy of a vb_graph_desc
  int64_t nv = 0, nP = 0;
  std::vector<int32_t> kind, tdim;
  std::vector<uint32_t> flags;
  std::vector<int64_t> index, off;
  std::vector<double> aux, data;
  struct Store {
    int32_t type, nvars, cpf;
    int64_t count;
    std::vector<int32_t> vars;
    std::vector<double> consts, aux;
    double loss_a, loss_k;
  };
  std::vector<Store> stores;
  void load(const vb_graph_desc* g);
};

struct SymbolicOptions {
  int maxLumpWidth = 96;     // scalar columns per supernode
  double relaxFill = 0.35;   // admissible fraction of explicit zeros when merging columns
  int ndLeafBlocks = 300;    // nested-dissection leaf size in camera blocks (<= 0: plain RCM chain)
};

// throws std::runtime_error on malformed graphs
void buildSymbolic(const HostGraph& g, const SymbolicOptions& opt, Symbolic* S);

} // namespace vb
