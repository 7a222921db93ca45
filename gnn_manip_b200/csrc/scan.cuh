// Exclusive prefix sum over int32 (three small launches, no host sync, graph-capturable).
#pragma once
#include "common.cuh"

namespace gm {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048 values per block

inline int scan_num_blocks(long long n) { return n <= 0 ? 1 : ceil_div(n, kScanTile); }
// workspace: one int per block (+1)
inline size_t scan_workspace_ints(long long n) { return (size_t)scan_num_blocks(n) + 1; }

#ifdef __CUDACC__
__device__ __forceinline__ int block_exclusive_scan_256(int v, int* total, int* smem /*>= 9 ints*/) {
  // warp inclusive scan
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) smem[warp] = x;
  __syncthreads();
  if (warp == 0) {
    int w = lane < (kScanThreads / 32) ? smem[lane] : 0;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += y;
    }
    if (lane < 8) smem[lane] = w;  // inclusive warp totals
  }
  __syncthreads();
  int warp_off = warp == 0 ? 0 : smem[warp - 1];
  *total = smem[7];
  __syncthreads();
  return warp_off + x - v;
}

// Phase A: per-block totals. `count_ptr` (optional) gives the live length on device (<= n).
static __global__ void scan_block_sums_kernel(const int* __restrict__ in, long long n, int* __restrict__ block_sums) {
  __shared__ int sm[9];
  long long base = (long long)blockIdx.x * kScanTile + (long long)threadIdx.x * kScanItems;
  int s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    long long k = base + i;
    if (k < n) s += in[k];
  }
  int total;
  block_exclusive_scan_256(s, &total, sm);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// Phase B: one block turns block totals into exclusive offsets (serial over chunks of 256).
static __global__ void scan_block_offsets_kernel(int* __restrict__ block_sums, int nblocks) {
  __shared__ int sm[9];
  __shared__ int carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < nblocks; base += kScanThreads) {
    int k = base + threadIdx.x;
    int v = k < nblocks ? block_sums[k] : 0;
    int total;
    int ex = block_exclusive_scan_256(v, &total, sm);
    int carry = carry_s;
    if (k < nblocks) block_sums[k] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) block_sums[nblocks] = carry_s;
}

// Phase C: out[k] = exclusive prefix, out[n] = grand total.
static __global__ void scan_apply_kernel(const int* __restrict__ in, long long n, const int* __restrict__ block_sums,
                                  int nblocks, int* __restrict__ out) {
  __shared__ int sm[9];
  long long base = (long long)blockIdx.x * kScanTile + (long long)threadIdx.x * kScanItems;
  int v[kScanItems];
  int s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    long long k = base + i;
    v[i] = k < n ? in[k] : 0;
    s += v[i];
  }
  int total;
  int ex = block_exclusive_scan_256(s, &total, sm) + block_sums[blockIdx.x];
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    long long k = base + i;
    if (k < n) out[k] = ex;
    ex += v[i];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = block_sums[nblocks];
}

// Small inputs: one block walks the array tile by tile (one launch instead of three).
constexpr long long kScanSingleBlockMax = 1 << 14;
static __global__ void scan_single_block_kernel(const int* __restrict__ in, long long n, int* __restrict__ out,
                                                int* __restrict__ total_out, const int* __restrict__ n_dev,
                                                int n_mul) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  // optional device-side bound: only the first *n_dev * n_mul entries are live (the rest of `in` is zero and the
  // rest of `out` is never read); out[n_live] receives the total
  if (n_dev != nullptr) n = min(n, (long long)(*n_dev) * n_mul);
  __shared__ int sm[9];
  int carry = 0;
  for (long long base0 = 0; base0 < n; base0 += kScanTile) {
    const long long base = base0 + (long long)threadIdx.x * kScanItems;
    int v[kScanItems];
    int s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
      const long long k = base + i;
      v[i] = k < n ? in[k] : 0;
      s += v[i];
    }
    int total;
    int ex = block_exclusive_scan_256(s, &total, sm) + carry;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
      const long long k = base + i;
      if (k < n) out[k] = ex;
      ex += v[i];
    }
    carry += total;
  }
  if (threadIdx.x == 0) {
    out[n] = carry;
    if (total_out) *total_out = carry;
  }
}

static __global__ void scan_copy_total_kernel(const int* __restrict__ out, long long n, int* __restrict__ total_out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *total_out = out[n];
}

// out has n+1 entries; `ws` has scan_workspace_ints(n) ints. in/out may alias only if identical pointers are
// NOT used (phase C reads `in` after phase A; aliasing in==out is safe because each block reads its own
// items before writing them and never reads another block's items).
// total_out (optional, device): receives the grand total as well.
inline int exclusive_scan_i32(const int* in, int* out, long long n, int* ws, cudaStream_t stream,
                              int* total_out = nullptr, const int* n_dev = nullptr, int n_mul = 1) {
  if (n <= kScanSingleBlockMax) {
    GM_CUDA(launch_pdl(scan_single_block_kernel, 1, kScanThreads, 0, stream, 1, in, n, out, total_out, n_dev, n_mul));
    GM_LAUNCHED(1);
    GM_LAUNCH_CHECK("scan_single_block_kernel");
    return GM_OK;
  }
  const int nb = scan_num_blocks(n);
  scan_block_sums_kernel<<<nb, kScanThreads, 0, stream>>>(in, n, ws);
  scan_block_offsets_kernel<<<1, kScanThreads, 0, stream>>>(ws, nb);
  scan_apply_kernel<<<nb, kScanThreads, 0, stream>>>(in, n, ws, nb, out);
  GM_LAUNCHED(3);
  if (total_out) {
    scan_copy_total_kernel<<<1, 32, 0, stream>>>(out, n, total_out);
    GM_LAUNCHED(1);
  }
  GM_LAUNCH_CHECK("exclusive_scan_i32");
  return GM_OK;
}
#endif

}  // namespace gm
