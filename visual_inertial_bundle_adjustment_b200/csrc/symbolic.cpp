// See symbolic.hpp.  Pure host C++17; no CUDA here.
#include "symbolic.hpp"
#include <algorithm>
#include <numeric>
#include <stdexcept>
#include <queue>
#include <cstring>
#include <chrono>
#include <cstdio>
#include <cstdlib>

namespace vb {

void HostGraph::load(const vb_graph_desc* g) {
  if (!g || g->num_vars <= 0 || !g->var_kind || !g->var_tdim || !g->var_index || !g->var_off || !g->var_data)
    throw std::runtime_error("vb_graph_desc: missing variable arrays");
  nv = g->num_vars;
  nP = g->num_points;
  kind.assign(g->var_kind, g->var_kind + nv);
  tdim.assign(g->var_tdim, g->var_tdim + nv);
  flags.assign(nv, 0u);
  if (g->var_flags) flags.assign(g->var_flags, g->var_flags + nv);
  index.assign(g->var_index, g->var_index + nv);
  off.assign(g->var_off, g->var_off + nv);
  aux.assign(nv, 0.0);
  if (g->var_aux) aux.assign(g->var_aux, g->var_aux + nv);
  data.assign(g->var_data, g->var_data + g->var_data_len);
  stores.resize(g->num_stores);
  for (int s = 0; s < g->num_stores; s++) {
    const vb_factor_store& fs = g->stores[s];
    Store& st = stores[s];
    st.type = fs.type; st.nvars = fs.nvars; st.cpf = fs.consts_per_factor; st.count = fs.count;
    if (fs.count > 0 && (!fs.vars || (fs.consts_per_factor > 0 && !fs.consts)))
      throw std::runtime_error("vb_factor_store: null table");
    st.vars.assign(fs.vars, fs.vars + fs.count * fs.nvars);
    for (int32_t v : st.vars)
      if (v < 0 || v >= nv) throw std::runtime_error("vb_factor_store: variable id out of range");
    st.consts.assign(fs.consts, fs.consts + fs.count * (int64_t)fs.consts_per_factor);
    if (fs.aux && fs.aux_len > 0) st.aux.assign(fs.aux, fs.aux + fs.aux_len);
    st.loss_a = fs.loss_a; st.loss_k = fs.loss_k;
  }
}

static void sortUnique(std::vector<int64_t>& v) {
  std::sort(v.begin(), v.end());
  v.erase(std::unique(v.begin(), v.end()), v.end());
}

int32_t Symbolic::rowInLump(int32_t L, int32_t c) const {
  if (camLump[c] == L) return camStart[c] - camStart[lumpFirst[L]];
  const int32_t* b = lumpRowCam.data() + lumpRowPtr[L];
  const int32_t* e = lumpRowCam.data() + lumpRowPtr[L + 1];
  const int32_t* p = std::lower_bound(b, e, c);
  if (p == e || *p != c) return -1;
  return lumpRowOff[p - lumpRowCam.data()];
}

bool Symbolic::block(int64_t ri, int64_t rj, BlockRef* out) const {
  int64_t pi = pos[ri], pj = pos[rj];
  int32_t tr = 0;
  if (pi < pj) { std::swap(pi, pj); tr = 1; }
  if (pj < nP) { // landmark column
    if (pi == pj) { *out = {lmOff[pj], 3, 0}; return true; }
    if (pi < nP) return false;
    const int32_t c = (int32_t)(pi - nP);
    const int32_t* b = lmRowCam.data() + lmRowPtr[pj];
    const int32_t* e = lmRowCam.data() + lmRowPtr[pj + 1];
    const int32_t* p = std::lower_bound(b, e, c);
    if (p == e || *p != c) return false;
    *out = {lmOff[pj] + (int64_t)lmRowOff[p - lmRowCam.data()] * 3, 3, tr};
    return true;
  }
  const int32_t cj = (int32_t)(pj - nP), ci = (int32_t)(pi - nP);
  const int32_t L = camLump[cj];
  const int32_t row = rowInLump(L, ci);
  if (row < 0) return false;
  const int32_t col = camStart[cj] - camStart[lumpFirst[L]];
  *out = {lumpOff[L] + (int64_t)row * lumpWidth[L] + col, lumpWidth[L], tr};
  return true;
}

void buildSymbolic(const HostGraph& g, const SymbolicOptions& opt, Symbolic* Sp) {
  Symbolic& S = *Sp;
  const bool verbose = std::getenv("VB_VERBOSE") != nullptr;
  auto tPrev = std::chrono::steady_clock::now();
  auto stage = [&](const char* what) {
    const auto now = std::chrono::steady_clock::now();
    if (verbose) std::fprintf(stderr, "[viba_b200] symbolic: %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - tPrev).count());
    tPrev = now;
  };
  const int64_t nP = g.nP;
  int64_t nB = 0;
  for (int64_t v = 0; v < g.nv; v++) if (g.index[v] >= 0) nB = std::max(nB, g.index[v] + 1);
  if (nP < 0 || nP > nB) throw std::runtime_error("num_points out of range");
  const int64_t nC = nB - nP;
  if (nC >= (1ll << 31) || nP >= (1ll << 31)) throw std::runtime_error("too many blocks");
  S.nB = nB; S.nP = nP; S.nC = nC;
  S.bsize.assign(nB, 0);
  for (int64_t v = 0; v < g.nv; v++)
    if (g.index[v] >= 0) {
      if (S.bsize[g.index[v]] != 0) throw std::runtime_error("duplicate registration index");
      if (g.tdim[v] <= 0 || g.tdim[v] > 23) throw std::runtime_error("tangent dim out of range");
      S.bsize[g.index[v]] = g.tdim[v];
    }
  for (int64_t b = 0; b < nB; b++) if (S.bsize[b] == 0) throw std::runtime_error("registration indices not contiguous");
  for (int64_t p = 0; p < nP; p++) if (S.bsize[p] != 3) throw std::runtime_error("elimination range must hold 3-dim landmarks");

  // ---- adjacency: landmark -> camera blocks, camera <-> camera (direct) --------------------
  // Both key lists are sorted and made unique below, so the two filters here only have to be safe (never
  // drop a key that was not pushed before): `camLast[c] == lm` means (lm, c) was pushed while this landmark
  // was being walked, `recent` is a direct-mapped cache of camera-pair keys.  On landmark-major factor lists
  // they remove most of the duplicates (one (landmark, calibration) pair per observation, one
  // (pose, calibration) pair per observation of a rig) before they reach the sort.
  std::vector<int32_t> pairLm, pairCam;   // (landmark, camera block) incidences, duplicates possible
  std::vector<int64_t> ccKeys;
  std::vector<int64_t> camLast(nC, -1);
  constexpr int kRecentBits = 16;
  std::vector<int64_t> recent(size_t(1) << kRecentBits, -1);
  for (const auto& st : g.stores) {
    for (int64_t k = 0; k < st.count; k++) {
      const int32_t* v = &st.vars[k * st.nvars];
      int64_t lm = -1;
      int64_t cams[16]; int nc = 0;
      for (int i = 0; i < st.nvars; i++) {
        const int64_t ii = g.index[v[i]];
        if (ii < 0) continue;
        if (ii < nP) { if (lm >= 0 && lm != ii) throw std::runtime_error("factor couples two landmarks"); lm = ii; }
        else cams[nc++] = ii - nP;
      }
      if (lm >= 0)
        for (int a = 0; a < nc; a++) {
          if (camLast[cams[a]] == lm) continue;
          camLast[cams[a]] = lm;
          pairLm.push_back((int32_t)lm);
          pairCam.push_back((int32_t)cams[a]);
        }
      for (int a = 0; a < nc; a++)
        for (int b = 0; b < a; b++) {
          if (cams[a] == cams[b]) continue;
          const int64_t key = std::max(cams[a], cams[b]) * nC + std::min(cams[a], cams[b]);
          int64_t& slot = recent[(size_t)((uint64_t)key * 0x9E3779B97F4A7C15ull >> (64 - kRecentBits))];
          if (slot == key) continue;
          slot = key;
          ccKeys.push_back(key);
        }
      if (ccKeys.size() > (size_t)(1 << 24)) sortUnique(ccKeys);
    }
  }
  stage("collect adjacency keys");
  sortUnique(ccKeys);
  // landmark -> cams CSR (camera registration order c = reg - nP), cameras ascending and unique per landmark:
  // bucket the incidences by landmark, then sort each (short) bucket
  std::vector<int64_t> lmPtr(nP + 1, 0);
  std::vector<int32_t> lmCam;
  {
    std::vector<int64_t> bucket(nP + 1, 0);
    for (int32_t p : pairLm) bucket[p + 1]++;
    for (int64_t p = 0; p < nP; p++) bucket[p + 1] += bucket[p];
    std::vector<int32_t> tmp(pairLm.size());
    {
      std::vector<int64_t> fill(bucket.begin(), bucket.end() - 1);
      for (size_t i = 0; i < pairLm.size(); i++) tmp[fill[pairLm[i]]++] = pairCam[i];
    }
    std::vector<int32_t>().swap(pairLm);
    std::vector<int32_t>().swap(pairCam);
    lmCam.reserve(tmp.size());
    for (int64_t p = 0; p < nP; p++) {
      int32_t* b = tmp.data() + bucket[p];
      int32_t* e = tmp.data() + bucket[p + 1];
      std::sort(b, e);
      e = std::unique(b, e);
      lmCam.insert(lmCam.end(), b, e);
      lmPtr[p + 1] = (int64_t)lmCam.size();
    }
  }
  stage("sort adjacency keys");
  // cam -> landmarks CSR
  std::vector<int64_t> clPtr(nC + 1, 0);
  for (int32_t c : lmCam) clPtr[c + 1]++;
  for (int64_t c = 0; c < nC; c++) clPtr[c + 1] += clPtr[c];
  std::vector<int32_t> clLm(lmCam.size());
  {
    std::vector<int64_t> fill(clPtr.begin(), clPtr.end() - 1);
    for (int64_t p = 0; p < nP; p++)
      for (int64_t k = lmPtr[p]; k < lmPtr[p + 1]; k++) clLm[fill[lmCam[k]]++] = (int32_t)p;
  }
  // cam <-> cam direct CSR (symmetric)
  std::vector<int64_t> ccPtr(nC + 1, 0);
  for (int64_t key : ccKeys) { ccPtr[key / nC + 1]++; ccPtr[key % nC + 1]++; }
  for (int64_t c = 0; c < nC; c++) ccPtr[c + 1] += ccPtr[c];
  std::vector<int32_t> ccAdj(ccPtr[nC]);
  {
    std::vector<int64_t> fill(ccPtr.begin(), ccPtr.end() - 1);
    for (int64_t key : ccKeys) {
      const int32_t a = (int32_t)(key / nC), b = (int32_t)(key % nC);
      ccAdj[fill[a]++] = b; ccAdj[fill[b]++] = a;
    }
  }

  stage("adjacency CSR");
  // ---- ordering of the camera side -----------------------------------------------------------
  // Nested dissection by BFS level structures on the Schur-filled graph (walked through the
  // landmark adjacency, never materialised): the visual-inertial graph is a chain in time, so a
  // BFS level is a time slab and makes a vertex separator; the two halves become independent
  // subtrees of the elimination tree (parallel on one GPU, the shard boundary across GPUs).
  // Leaves and separators are ordered by (reverse) Cuthill-McKee.  opt.ndLeafBlocks <= 0 gives
  // plain RCM (a single chain).
  std::vector<int32_t> orderC; orderC.reserve(nC);
  {
    std::vector<int32_t> deg(nC);
    for (int64_t c = 0; c < nC; c++) deg[c] = (int32_t)((ccPtr[c + 1] - ccPtr[c]) + (clPtr[c + 1] - clPtr[c]));
    std::vector<int32_t> inSet(nC, -1), seen(nC, -1), lmSeen(nP, -1);
    int32_t bfsStamp = 0, setStamp = 0;
    // BFS inside the node set stamped `set`; appends discovery order to out, level starts to lv
    auto bfs = [&](int32_t root, int32_t set, std::vector<int32_t>& out, std::vector<int32_t>* lv) {
      const int32_t stamp = ++bfsStamp;
      size_t head = out.size(), levelEnd;
      out.push_back(root); seen[root] = stamp;
      std::vector<int32_t> nb;
      if (lv) lv->push_back((int32_t)head);
      levelEnd = out.size();
      while (head < out.size()) {
        if (head == levelEnd) { if (lv) lv->push_back((int32_t)head); levelEnd = out.size(); }
        const int32_t u = out[head++];
        nb.clear();
        for (int64_t k = ccPtr[u]; k < ccPtr[u + 1]; k++) {
          const int32_t w = ccAdj[k];
          if (inSet[w] == set && seen[w] != stamp) { seen[w] = stamp; nb.push_back(w); }
        }
        for (int64_t k = clPtr[u]; k < clPtr[u + 1]; k++) {
          const int32_t p = clLm[k];
          if (lmSeen[p] == stamp) continue;
          lmSeen[p] = stamp;
          for (int64_t q = lmPtr[p]; q < lmPtr[p + 1]; q++) {
            const int32_t w = lmCam[q];
            if (inSet[w] == set && seen[w] != stamp) { seen[w] = stamp; nb.push_back(w); }
          }
        }
        std::sort(nb.begin(), nb.end(), [&](int32_t a, int32_t b) { return deg[a] != deg[b] ? deg[a] < deg[b] : a < b; });
        out.insert(out.end(), nb.begin(), nb.end());
      }
      if (lv) lv->push_back((int32_t)out.size());
    };
    auto weight = [&](int32_t c) { return (int64_t)S.bsize[nP + c]; };
    struct Job { std::vector<int32_t> nodes; bool dissect; };
    // explicit stack; emitted segments are appended to orderC in elimination order
    std::vector<Job> stack;
    {
      Job all; all.nodes.resize(nC); std::iota(all.nodes.begin(), all.nodes.end(), 0); all.dissect = true;
      stack.push_back(std::move(all));
    }
    while (!stack.empty()) {
      Job job = std::move(stack.back());
      stack.pop_back();
      if (job.nodes.empty()) continue;
      const int32_t set = ++setStamp;
      for (int32_t c : job.nodes) inSet[c] = set;
      // min-degree start, then one sweep to a pseudo-peripheral root
      int32_t start = job.nodes[0];
      for (int32_t c : job.nodes) if (deg[c] < deg[start] || (deg[c] == deg[start] && c < start)) start = c;
      std::vector<int32_t> comp, lv;
      bfs(start, set, comp, nullptr);
      const int32_t root = comp.back();
      comp.clear();
      bfs(root, set, comp, &lv);
      if (comp.size() < job.nodes.size()) { // disconnected: the rest is an independent job (no separator)
        Job rest; rest.dissect = job.dissect;
        for (int32_t c : job.nodes) if (seen[c] != bfsStamp) rest.nodes.push_back(c);
        stack.push_back(std::move(rest));
      }
      const int nlev = (int)lv.size() - 1;
      if (!job.dissect || opt.ndLeafBlocks <= 0 || (int64_t)comp.size() <= opt.ndLeafBlocks || nlev < 5) {
        orderC.insert(orderC.end(), comp.rbegin(), comp.rend()); // reverse Cuthill-McKee
        continue;
      }
      // separator = the lightest level among those keeping both sides within 35..65 % of the weight
      std::vector<int64_t> cum(nlev + 1, 0);
      for (int l = 0; l < nlev; l++) { int64_t w = 0; for (int32_t i = lv[l]; i < lv[l + 1]; i++) w += weight(comp[i]); cum[l + 1] = cum[l] + w; }
      int best = -1;
      for (int l = 1; l + 1 < nlev; l++) {
        const double left = (double)cum[l] / cum[nlev], right = (double)(cum[nlev] - cum[l + 1]) / cum[nlev];
        if (left < 0.3 || right < 0.3) continue;
        if (best < 0 || cum[l + 1] - cum[l] < cum[best + 1] - cum[best]) best = l;
      }
      if (best < 0) best = nlev / 2;
      Job A, B, Sep;
      A.dissect = B.dissect = true; Sep.dissect = false;
      A.nodes.assign(comp.begin(), comp.begin() + lv[best]);
      Sep.nodes.assign(comp.begin() + lv[best], comp.begin() + lv[best + 1]);
      B.nodes.assign(comp.begin() + lv[best + 1], comp.end());
      // LIFO: A is processed (and emitted) first, then B, then the separator
      stack.push_back(std::move(Sep));
      stack.push_back(std::move(B));
      stack.push_back(std::move(A));
    }
    if ((int64_t)orderC.size() != nC) throw std::runtime_error("ordering lost camera blocks");
  }
  stage("nested dissection ordering");
  std::vector<int32_t> cpos(nC); // camera registration c -> camera position
  for (int64_t i = 0; i < nC; i++) cpos[orderC[i]] = (int32_t)i;

  S.pos.assign(nB, 0); S.posToReg.assign(nB, 0);
  for (int64_t p = 0; p < nP; p++) { S.pos[p] = p; S.posToReg[p] = p; }
  for (int64_t c = 0; c < nC; c++) { S.pos[nP + c] = nP + cpos[c]; S.posToReg[nP + cpos[c]] = nP + c; }
  S.vecStart.assign(nB + 1, 0);
  for (int64_t q = 0; q < nB; q++) S.vecStart[q + 1] = S.vecStart[q] + S.bsize[S.posToReg[q]];
  S.order = S.vecStart[nB];
  S.camStart.assign(nC + 1, 0);
  for (int64_t q = 0; q < nC; q++) S.camStart[q + 1] = S.camStart[q] + S.bsize[S.posToReg[nP + q]];

  // ---- landmark panels ------------------------------------------------------------------------
  S.lmOff.assign(nP + 1, 0); S.lmRowPtr = lmPtr;
  S.lmRowCam.resize(lmCam.size()); S.lmRowOff.resize(lmCam.size());
  S.schurFlops = 0;
  {
    int64_t off = 0;
    for (int64_t p = 0; p < nP; p++) {
      S.lmOff[p] = off;
      const int64_t b = lmPtr[p], e = lmPtr[p + 1];
      for (int64_t k = b; k < e; k++) S.lmRowCam[k] = cpos[lmCam[k]];
      std::sort(S.lmRowCam.begin() + b, S.lmRowCam.begin() + e);
      int32_t row = 3;
      for (int64_t k = b; k < e; k++) { S.lmRowOff[k] = row; row += S.camStart[S.lmRowCam[k] + 1] - S.camStart[S.lmRowCam[k]]; }
      off += (int64_t)row * 3;
      const double m = row - 3;
      S.schurFlops += 3.0 * m * m + 9.0 * m + 9.0;
    }
    S.lmOff[nP] = off;
    S.lmDataSize = off;
  }

  stage("landmark panels");
  // ---- symbolic Cholesky of the reduced camera system (column structures via the etree) ------
  std::vector<std::vector<int32_t>> cstruct(nC);
  std::vector<int32_t> parent(nC, -1);
  {
    // landmarks bucketed by their first (minimum-position) neighbour
    std::vector<int64_t> byMinPtr(nC + 1, 0);
    for (int64_t p = 0; p < nP; p++) if (lmPtr[p + 1] > lmPtr[p]) byMinPtr[S.lmRowCam[lmPtr[p]] + 1]++;
    for (int64_t c = 0; c < nC; c++) byMinPtr[c + 1] += byMinPtr[c];
    std::vector<int32_t> byMin(byMinPtr[nC]);
    {
      std::vector<int64_t> fill(byMinPtr.begin(), byMinPtr.end() - 1);
      for (int64_t p = 0; p < nP; p++) if (lmPtr[p + 1] > lmPtr[p]) byMin[fill[S.lmRowCam[lmPtr[p]]]++] = (int32_t)p;
    }
    std::vector<std::vector<int32_t>> children(nC);
    std::vector<int32_t> mark(nC, -1);
    for (int32_t j = 0; j < (int32_t)nC; j++) {
      std::vector<int32_t>& st = cstruct[j];
      auto add = [&](int32_t w) { if (w > j && mark[w] != j) { mark[w] = j; st.push_back(w); } };
      const int32_t cj = orderC[j];
      for (int64_t k = ccPtr[cj]; k < ccPtr[cj + 1]; k++) add(cpos[ccAdj[k]]);
      for (int64_t k = byMinPtr[j]; k < byMinPtr[j + 1]; k++) {
        const int32_t p = byMin[k];
        for (int64_t q = lmPtr[p] + 1; q < lmPtr[p + 1]; q++) add(S.lmRowCam[q]);
      }
      for (int32_t c : children[j]) {
        for (int32_t w : cstruct[c]) add(w);
      }
      std::vector<int32_t>().swap(children[j]);
      std::sort(st.begin(), st.end());
      if (!st.empty()) { parent[j] = st[0]; children[st[0]].push_back(j); }
    }
  }

  stage("column structures");
  // ---- supernodes ("lumps"): chains of columns merged under a width cap and a fill budget ----
  auto csize = [&](int32_t c) { return S.camStart[c + 1] - S.camStart[c]; };
  S.camLump.assign(nC, 0);
  S.lumpFirst.clear(); S.lumpWidth.clear(); S.lumpRows.clear(); S.lumpOff.clear();
  S.lumpRowPtr.assign(1, 0); S.lumpRowCam.clear(); S.lumpRowOff.clear();
  {
    int32_t j = 0;
    std::vector<int32_t> cur, merged;
    while (j < (int32_t)nC) {
      const int32_t j0 = j;
      cur = cstruct[j];
      int32_t width = csize(j);
      auto scal = [&](const std::vector<int32_t>& v) { int64_t s = 0; for (int32_t c : v) s += csize(c); return s; };
      double exact = (double)csize(j) * (csize(j) + scal(cur));
      while (j + 1 < (int32_t)nC && parent[j] == j + 1 && width + csize(j + 1) <= opt.maxLumpWidth) {
        const int32_t n = j + 1;
        merged.clear();
        // (cur \ {n}) U struct(n)
        std::set_union(cur.begin() + (cur.size() && cur[0] == n ? 1 : 0), cur.end(), cstruct[n].begin(), cstruct[n].end(), std::back_inserter(merged));
        const double exactN = exact + (double)csize(n) * (csize(n) + scal(cstruct[n]));
        // lower-trapezoid area of the merged panel
        const int64_t rowsBelow = scal(merged);
        double area = 0;
        int32_t rem = width + csize(n);
        for (int32_t c = j0; c <= n; c++) { area += (double)csize(c) * (rem + rowsBelow); rem -= csize(c); }
        if (area - exactN > opt.relaxFill * area) break;
        cur.swap(merged);
        exact = exactN;
        width += csize(n);
        j = n;
      }
      const int32_t L = (int32_t)S.lumpFirst.size();
      S.lumpFirst.push_back(j0);
      for (int32_t c = j0; c <= j; c++) S.camLump[c] = L;
      S.lumpWidth.push_back(width);
      int32_t row = width;
      for (int32_t c : cur) { S.lumpRowCam.push_back(c); S.lumpRowOff.push_back(row); row += csize(c); }
      S.lumpRowPtr.push_back((int64_t)S.lumpRowCam.size());
      S.lumpRows.push_back(row);
      j++;
    }
    S.nL = (int64_t)S.lumpFirst.size();
    S.lumpFirst.push_back((int32_t)nC);
  }
  std::vector<std::vector<int32_t>>().swap(cstruct);

  // ---- storage offsets, etree over lumps, levels, work estimates ------------------------------
  S.lumpOff.assign(S.nL, 0);
  S.lumpParent.assign(S.nL, -1);
  S.lumpLevel.assign(S.nL, 0);
  S.cholFlops = 0; S.maxLumpWidth = 0; S.maxLumpRows = 0;
  int64_t off = (S.lmDataSize + 1) & ~int64_t(1); // panels start 16-byte aligned (16-byte async copies)
  for (int64_t L = 0; L < S.nL; L++) {
    S.lumpOff[L] = off;
    off += ((int64_t)S.lumpRows[L] * S.lumpWidth[L] + 1) & ~int64_t(1);
    const double w = S.lumpWidth[L], r = S.lumpRows[L] - S.lumpWidth[L];
    S.cholFlops += w * w * w / 3.0 + w * w * r + w * r * r;
    S.maxLumpWidth = std::max(S.maxLumpWidth, S.lumpWidth[L]);
    S.maxLumpRows = std::max(S.maxLumpRows, S.lumpRows[L]);
    if (S.lumpRowPtr[L + 1] > S.lumpRowPtr[L]) S.lumpParent[L] = S.camLump[S.lumpRowCam[S.lumpRowPtr[L]]];
  }
  S.dataSize = off;
  int32_t maxLevel = 0;
  for (int64_t L = 0; L < S.nL; L++) {
    const int32_t p = S.lumpParent[L];
    if (p >= 0) S.lumpLevel[p] = std::max(S.lumpLevel[p], S.lumpLevel[L] + 1);
    maxLevel = std::max(maxLevel, S.lumpLevel[L]);
  }
  S.levels.assign(S.nL ? maxLevel + 1 : 0, {});
  for (int64_t L = 0; L < S.nL; L++) S.levels[S.lumpLevel[L]].push_back((int32_t)L);
  stage("supernodes and levels");
}

} // namespace vb
