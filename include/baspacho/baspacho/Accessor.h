// baspacho/baspacho/Accessor.h — facade of BaSpaCho's PermutedCoalescedAccessor as the reference's direct path uses it
// (SURVEY.md §2.3): registration index -> permuted scalar offset (paramStart), block size (paramSize), and writable
// views of the Hessian blocks inside the flat `hess` array (block, diagBlock) —
//   lib/small_thing/Factor.h:612,633,643 (gradient segment, off-diagonal and diagonal block accumulation),
//   lib/small_thing/Variable.h:360 (step segment of a variable), lib/small_thing/Optimizer.cpp:142 (damping).
// Here the layout is this library's (landmark panels + supernode panels), read through
// vb_get_param_start / vb_block_offset.  Eigen is not available in this repo, so a block is a small strided view
// instead of an Eigen::Map; `transposed` says the stored block is (j, i) and the view flips the indices, as BaSpaCho's
// accessor does for the lower triangle.
#pragma once
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../viba_b200.h"
#include "DebugMacros.h"

namespace BaSpaCho {

template <typename T>
struct BlockView {   // rows x cols view, element (r, c) at data[r * rowStride + c * colStride]
  T* data = nullptr;
  int64_t rows = 0, cols = 0, rowStride = 0, colStride = 0;
  T& operator()(int64_t r, int64_t c) const { return data[r * rowStride + c * colStride]; }
};

class PermutedCoalescedAccessor {
 public:
  PermutedCoalescedAccessor() = default;
  explicit PermutedCoalescedAccessor(const vb_problem* p) : p_(p) {
    const int64_t nb = vb_num_blocks(p);
    start_.resize((size_t)nb);
    if (vb_get_param_start(p, start_.data(), nb) != VB_OK) throw std::runtime_error(std::string("BaSpaCho facade: ") + vb_last_error());
    // paramStart is in permuted order: sizes come from the sorted starts
    std::vector<int64_t> sorted(start_.begin(), start_.end());
    sorted.push_back(vb_order(p));
    std::sort(sorted.begin(), sorted.end());
    size_.resize((size_t)nb);
    for (int64_t i = 0; i < nb; i++) {
      const auto it = std::upper_bound(sorted.begin(), sorted.end(), start_[(size_t)i]);
      size_[(size_t)i] = *it - start_[(size_t)i];
    }
  }
  int64_t paramStart(int64_t i) const { return start_[(size_t)i]; }   // scalar offset in the permuted vectors
  int64_t paramSize(int64_t i) const { return size_[(size_t)i]; }
  // block (i, j) of the symmetric Hessian, i != j, as a paramSize(i) x paramSize(j) view
  template <typename T>
  BlockView<T> block(T* data, int64_t i, int64_t j) const {
    int64_t off = 0, stride = 0;
    int32_t tr = 0;
    if (vb_block_offset(p_, i, j, &off, &stride, &tr) != VB_OK) throw std::runtime_error(std::string("BaSpaCho facade: ") + vb_last_error());
    BlockView<T> v;
    v.data = data + off; v.rows = paramSize(i); v.cols = paramSize(j);
    if (tr) { v.rowStride = 1; v.colStride = stride; } else { v.rowStride = stride; v.colStride = 1; }
    return v;
  }
  template <typename T>
  BlockView<T> diagBlock(T* data, int64_t i) const { return block(data, i, i); }

 private:
  const vb_problem* p_ = nullptr;
  std::vector<int64_t> start_, size_;
};

}  // namespace BaSpaCho
