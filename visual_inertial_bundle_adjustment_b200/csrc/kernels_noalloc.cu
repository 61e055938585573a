// vb_dot_no_alloc: a.b over caller-owned device vectors WITHOUT any allocation — for callers that cannot reach cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the
// reference.
//
// It is the allocation-free variant of the dot-product kernel behind vb_dot / vb_step_stats (vecOpKernel op 2 in
// kernels_factors.cu), whose accumulator is scratch that the problem cudaMalloc's.  Here the scratch — one partial sum
// per CTA plus the "last CTA done" ticket — lives in a file-scope __device__ array of fixed size:
//   kNaMaxBlocks partial sums (16 384 doubles = 128 KiB) + one 256-byte line for the ticket,
// held by every GPU that loads the library until the process exits.  One CTA covers kNaChunk = 16 384 elements, so the
// array bounds the vector length at VB_DOT_NO_ALLOC_MAX_N = 16 384 x 16 384 = 2^28 elements (2 GiB per vector).
//
// In host code `g_naScratch` is a host-side stand-in, not a device address: the address every launch passes comes from
// cudaGetSymbolAddress, taken once per device, with that device current (scratchFor below).
//
// The result is deterministic (fixed summation order: 256 strided lanes per CTA, shuffle tree, CTA partials in index
// order), unlike the fp64-atomic accumulation of vb_dot.  All calls on one GPU share the array: two calls whose kernels
// overlap on the GPU (different streams, no dependency between them) race on it; calls on one stream, or ordered by
// events, are safe.  The ticket is the only cell that must start at zero: it is cleared on the stream before every launch
// (and by the last CTA after it); the partial sums are fully overwritten before they are read.
#include <mutex>
#include "vb_internal.hpp"
#include "../../include/viba_b200.h"

namespace vb {

void setLastError(const std::string& text);   // api.cu (thread-local text behind vb_last_error)

constexpr int kNaThreads = 256;
constexpr int kNaChunk = 16384;        // elements per CTA
constexpr int kNaMaxBlocks = 16384;    // partial sums the scratch holds
constexpr int kNaTicketAt = kNaMaxBlocks;          // the ticket's 256-byte line starts here
constexpr int kNaScratchDoubles = kNaMaxBlocks + 32;
static_assert((int64_t)kNaChunk * kNaMaxBlocks == VB_DOT_NO_ALLOC_MAX_N, "header constant out of step with the scratch size");

static __device__ __align__(256) double g_naScratch[kNaScratchDoubles];

// sum over the CTA in a fixed order; the result is valid in thread 0
__device__ double naBlockSum(double v, double* sh) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < kNaThreads / 32; w++) s += sh[w];
  __syncthreads();
  return s;
}

__global__ void __launch_bounds__(kNaThreads) dotNoAllocKernel(const double* __restrict__ a, const double* __restrict__ b, int64_t n,
                                                               double* scratch, double* out) {
  __shared__ double sh[kNaThreads / 32];
  __shared__ int isLast;
  const int64_t beg = (int64_t)blockIdx.x * kNaChunk;
  const int64_t end = min(n, beg + (int64_t)kNaChunk);
  double acc = 0.0;
  for (int64_t i = beg + threadIdx.x; i < end; i += kNaThreads) acc = fma(a[i], b[i], acc);
  const double part = naBlockSum(acc, sh);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + kNaTicketAt);
  if (threadIdx.x == 0) {
    __stcg(scratch + blockIdx.x, part);
    __threadfence();                                  // the partial is visible before the ticket is taken
    isLast = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!isLast) return;
  __threadfence();
  double s = 0.0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += kNaThreads) s += __ldcg(scratch + i);
  const double total = naBlockSum(s, sh);
  if (threadIdx.x == 0) { out[0] = total; *ticket = 0u; }
}

// device address of the scratch array on `device` (which must be current): cudaGetSymbolAddress once per device
static double* scratchFor(int device) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static double* addr[kMaxDevices] = {};
  if (device < 0 || device >= kMaxDevices) throw std::invalid_argument("vb_dot_no_alloc: device index out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    void* ptr = nullptr;
    VB_CUDA(cudaGetSymbolAddress(&ptr, g_naScratch));
    if (!ptr) throw std::runtime_error("cudaGetSymbolAddress returned a null address (CUDA)");
    addr[device] = static_cast<double*>(ptr);
  }
  return addr[device];
}

}  // namespace vb

extern "C" int vb_dot_no_alloc(int device, const double* a_dev, const double* b_dev, int64_t n, double* out_dev, void* cuda_stream) {
  using namespace vb;
  try {
    if (n < 0) throw std::invalid_argument("vb_dot_no_alloc: negative length");
    if (n > VB_DOT_NO_ALLOC_MAX_N)
      throw std::invalid_argument("vb_dot_no_alloc: n = " + std::to_string(n) + " exceeds VB_DOT_NO_ALLOC_MAX_N = " +
                                  std::to_string((long long)VB_DOT_NO_ALLOC_MAX_N) + " (the fixed scratch holds one partial sum per 16384 elements)");
    if (!out_dev || (n > 0 && (!a_dev || !b_dev))) throw std::invalid_argument("vb_dot_no_alloc: null device pointer");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) throw std::runtime_error("no CUDA device (vb_dot_no_alloc has no CPU fallback)");
    if (device < 0 || device >= count) throw std::invalid_argument("vb_dot_no_alloc: no such device");
    VB_CUDA(cudaSetDevice(device));
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    if (n == 0) { VB_CUDA(cudaMemsetAsync(out_dev, 0, sizeof(double), s)); return VB_OK; }
    double* scratch = scratchFor(device);
    VB_CUDA(cudaMemsetAsync(scratch + kNaTicketAt, 0, 256, s));   // the one cell that must start at zero
    const unsigned blocks = (unsigned)((n + kNaChunk - 1) / kNaChunk);
    dotNoAllocKernel<<<blocks, kNaThreads, 0, s>>>(a_dev, b_dev, n, scratch, out_dev);
    VB_CUDA(cudaGetLastError());
    return VB_OK;
  } catch (const std::invalid_argument& ex) { setLastError(ex.what()); return VB_ERR_ARG; }
  catch (const std::exception& ex) { setLastError(ex.what()); return VB_ERR_CUDA; }
}
