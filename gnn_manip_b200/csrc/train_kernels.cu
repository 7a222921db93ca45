// Rows E(2) + N3 -- kernels of the training step (examples/train_dyn.py:49-72: forward, L1 loss, backward, Adam).
//
// The reference trains through torch autograd over the same Encode-Process-Decode stack (gnn_manip/models/epd_gnn.py,
// deps/torch-graphnet/torch_graphnet/graph_networks.py); here every 128-wide Linear of forward, dgrad and wgrad is a
// tcgen05 GEMM with the bf16x3 operand split (fp32-equivalent products, fp32 accumulation in TMEM):
//   tr_gemm   Y = epilogue(X W^T)        rows x 128 x 128, operands in shared memory (K-major, SWIZZLE_128B), weights
//                                        resident per CTA; forward layers and dgrad (with the transposed pack)
//   tr_wgrad  dW += A^T B                128 x 128 x rows: the SAME shared-memory image of a row tile, read through
//                                        MN-major descriptors (features contiguous, rows along K), accumulated in
//                                        TMEM over all row tiles of a CTA; deterministic two-stage reduction
// LayerNorm forward / backward, the 4 / 25 / 3-wide first and last layers, the loss and Adam run on CUDA cores.
// Activations are stored fp32; every global access of the GEMM kernels is a whole 512 B row.
#include <mutex>
#include "train.cuh"

namespace gm {

namespace {

constexpr int kT = 128;                       // rows per tile
constexpr int kNT = 256;                      // threads: (lane quadrant q = warp & 3) x (column half hf = warp >> 2)
constexpr int kImage = 32768;                 // one 128 x 128 bf16 image
constexpr int kKBlock = 16384;                // 128 rows x 128 B: one 64-column K block
constexpr uint32_t kIdescKK = umma_idesc_bf16(kT, kTrH);                              // A, B K-major
constexpr uint32_t kIdescMN = umma_idesc_bf16(kT, kTrH) | (1u << 15) | (1u << 16);    // A, B MN-major (wgrad)

// byte offset of element (row, col) inside a SWIZZLE_128B image of 128-byte rows, 64 columns per block
__device__ __forceinline__ uint32_t img_off(int row, int col, int block_bytes) {
  const int kb = col >> 6, chunk = (col & 63) >> 3;
  return kb * block_bytes + row * 128 + ((chunk ^ (row & 7)) << 4) + ((col & 7) << 1);
}
// fp32 staging tile [128][128]: 16-byte chunk index XOR (row & 7)
__device__ __forceinline__ uint32_t stg_off(int row, int col) {
  return (uint32_t)row * 512u + ((((uint32_t)col >> 2) ^ ((uint32_t)row & 7u)) << 4) + (((uint32_t)col & 3u) << 2);
}

struct GemmSmem {
  uint64_t bar_w, bar_mma[2];
  uint32_t tmem_base;
  uint32_t pad;
  alignas(16) float bias[kTrH];
};

__device__ __forceinline__ void issue_gemm_kk(uint32_t a_s, uint32_t w_s, uint32_t tmem_d) {
  const uint64_t a0 = umma_desc_sw128(a_s), w0 = umma_desc_sw128(w_s);   // base descriptors once, one add per MMA operand
#pragma unroll
  for (int kb = 0; kb < 2; ++kb) {
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      const uint32_t off = kb * kKBlock + ks * 32;
      const uint64_t ah = umma_desc_advance(a0, off), al = umma_desc_advance(a0, kImage + off);
      const uint64_t wh = umma_desc_advance(w0, off), wl = umma_desc_advance(w0, kImage + off);
      umma_bf16_ss(tmem_d, ah, wh, kIdescKK, (kb | ks) ? 1u : 0u);
      umma_bf16_ss(tmem_d, ah, wl, kIdescKK, 1u);
      umma_bf16_ss(tmem_d, al, wh, kIdescKK, 1u);
    }
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void prefetch_l2_line(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Two tile slots per CTA (256 threads each, named barriers): the weights are shared, every slot has its own 64 KB
// buffer -- first the bf16 hi / lo operand images, then, once the MMAs have read them, the fp32 staging tile of the
// same rows -- and its own 128 accumulator columns.  The slots run the same serial tile loop out of phase, so that one
// slot's global loads / stores overlap the other's conversion, MMAs and epilogue.
constexpr int kGemmSlots = 2;
constexpr int kGemmThreads = kGemmSlots * kNT;

// kLN: the epilogue is LayerNorm (+ residual) instead of adds / ReLU / mask: with a whole row per warp the statistics are
// two warp reductions (same arithmetic as tr_ln_fwd_kernel), and the pre-normalisation rows never go to memory.
// kMode specialises the row epilogue at compile time (the generic version carries the predicated code of every option):
// 0 = any combination (run-time flags), 1 = neither adds, mask nor accumulate, 2 = mask only, 3 = accumulate only.
template <bool kLN, int kMode>
__global__ void __launch_bounds__(kGemmThreads, 1) tr_gemm_kernel(TrGemmArgs a) {
  const bool has_add1 = kMode == 0 && a.add1 != nullptr, has_add2 = kMode == 0 && a.add2 != nullptr;
  const bool has_mask = kMode == 2 || (kMode == 0 && a.mask != nullptr);
  const bool has_acc = kMode == 3 || (kMode == 0 && a.accumulate != 0);
  extern __shared__ char smem_raw[];
  char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  char* W = base;                              // weight images (hi, lo), resident
  GemmSmem* sh = reinterpret_cast<GemmSmem*>(base + (1 + kGemmSlots) * kTrChunkBytes);
  // (slot through a shuffle: provably warp-uniform, so the tile loop is uniform control flow for the compiler)
  const int slot = __shfl_sync(0xffffffffu, threadIdx.x / kNT, 0), tid = threadIdx.x % kNT;
  const int lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), q = warp & 3, hf = warp >> 2, row = q * 32 + lane;
  char* A = base + (1 + slot) * kTrChunkBytes; // operand images (hi, lo) of this slot ...
  char* T = A;                                 // ... and, after the MMAs, its fp32 staging tile
  const uint32_t bar_w = smem_u32(&sh->bar_w), bar_mma = smem_u32(&sh->bar_mma[slot]);
  if (threadIdx.x == 0) {
    mbar_init(bar_w, 1);
    mbar_init(smem_u32(&sh->bar_mma[0]), 1);
    mbar_init(smem_u32(&sh->bar_mma[1]), 1);
    fence_mbar_init();
    mbar_expect_tx(bar_w, kTrChunkBytes);
    for (int i = 0; i < 4; ++i) bulk_g2s(smem_u32(W) + i * kKBlock, (const char*)a.chunk + (size_t)i * kKBlock, kKBlock, bar_w);
  }
  if (threadIdx.x < 32) {
    tmem_alloc(smem_u32(&sh->tmem_base), 128 * kGemmSlots);
    tmem_relinquish();
  }
  for (int i = threadIdx.x; i < kTrH; i += kGemmThreads) sh->bias[i] = a.bias ? a.bias[i] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_base + 128u * slot;
  const int M = a.M_dev ? min(*a.M_dev, a.M) : a.M;
  const int n_tiles = (M + kT - 1) / kT;
  constexpr int R = kT / (kNT / 32);           // 16 rows per warp in the row-wise phases
  uint32_t ph = 0;
  bool w_ready = false;
  for (int tile = blockIdx.x + slot * gridDim.x; tile < n_tiles; tile += kGemmSlots * gridDim.x) {
    const int row0 = tile * kT, rows_valid = min(kT, M - row0);
    // ---- X rows -> bf16 hi / lo operand images (warp per row, coalesced 512 B) ----
#pragma unroll
    for (int i0 = 0; i0 < R; i0 += 8) {
      float4 x[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        // branch-free: rows past the end re-read the tile's last valid row and are zeroed afterwards
        const int r = warp * R + i0 + i;
        size_t src = (size_t)(row0 + min(r, rows_valid - 1));
        if (a.in_idx) src = (size_t)a.in_idx[src];
        x[i] = __ldg(reinterpret_cast<const float4*>(a.x + src * kTrH) + lane);
        if (r >= rows_valid) x[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = warp * R + i0 + i;
        uint2 hi, lo;
        split_bf16x2(x[i].x, x[i].y, hi.x, lo.x);
        split_bf16x2(x[i].z, x[i].w, hi.y, lo.y);
        const uint32_t o = img_off(r, 4 * lane, kKBlock);
        *reinterpret_cast<uint2*>(A + o) = hi;
        *reinterpret_cast<uint2*>(A + kImage + o) = lo;
      }
    }
    // The tile loop is a chain of dependent phases, so its global reads are requested well ahead: the X rows of this
    // slot's NEXT tile and the mask / accumulate rows of this tile's epilogue are pulled into L2 now (lane l: 128-byte
    // line l & 3 of row l >> 2, sixteen rows per warp in two steps).
    {
      const int nt = tile + kGemmSlots * (int)gridDim.x;
#pragma unroll
      for (int k2 = 0; k2 < 2; ++k2) {
        const int r = warp * R + 8 * k2 + (lane >> 2);
        const size_t off = (size_t)r * kTrH + 32 * (lane & 3);
        if (nt < n_tiles && a.in_idx == nullptr && nt * kT + r < M) prefetch_l2_line(a.x + (size_t)nt * kT * kTrH + off);
        if (r < rows_valid) {
          if (!kLN && has_mask) prefetch_l2_line(a.mask + (size_t)row0 * kTrH + off);
          if (!kLN && has_acc) prefetch_l2_line(a.out + (size_t)row0 * kTrH + off);
          if (kLN && a.res_out) prefetch_l2_line(a.res_in + (size_t)row0 * kTrH + off);
        }
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    named_bar_sync(1 + slot, kNT);
    if (tid == 0) {
      if (!w_ready) mbar_wait(bar_w, 0);
      tc_fence_after();
      issue_gemm_kk(smem_u32(A), smem_u32(W), tmem);
      umma_commit(bar_mma);
    }
    w_ready = true;
    __syncwarp();
    // source rows of the gathered adds for this warp's 16 rows (lane i < 16: row i), fetched under the MMAs
    int my_i1 = 0, my_i2 = 0;
    if constexpr (!kLN && kMode == 0) {
      const int rr = warp * R + (lane & 15);
      const bool rv = rr < rows_valid;
      my_i1 = (a.idx1 && rv) ? __ldg(a.idx1 + row0 + rr) : row0 + rr;
      my_i2 = (a.idx2 && rv) ? __ldg(a.idx2 + row0 + rr) : row0 + rr;
      if (!rv) my_i1 = my_i2 = 0;
    }
    mbar_wait(bar_mma, ph);                    // the MMAs have read the operand images: the buffer becomes the staging tile
    ph ^= 1u;
    tc_fence_after();
    // ---- accumulator (+ bias) -> staging tile: thread = (row, column half) ----
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int col0 = hf * 64 + h * 32;
      float v[32];
      tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)col0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 bs = *reinterpret_cast<const float4*>(sh->bias + col0 + 4 * j);
        *reinterpret_cast<float4*>(T + stg_off(row, col0 + 4 * j)) =
            make_float4(v[4 * j] + bs.x, v[4 * j + 1] + bs.y, v[4 * j + 2] + bs.z, v[4 * j + 3] + bs.w);
      }
    }
    tc_fence_before();
    named_bar_sync(1 + slot, kNT);
    // ---- row-wise epilogue: every global access is one coalesced 512 B row; four rows per step with all their
    //      loads (gathered adds, mask, accumulate target) issued before the first use ----
    if constexpr (kLN) {
      const float4 gg = reinterpret_cast<const float4*>(a.ln_g)[lane], bb = reinterpret_cast<const float4*>(a.ln_b)[lane];
#pragma unroll 1
      for (int i0 = 0; i0 < R; i0 += 4) {
        float4 ee[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = warp * R + i0 + i;
          ee[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (a.res_out && r < rows_valid) ee[i] = __ldg(reinterpret_cast<const float4*>(a.res_in + (size_t)(row0 + r) * kTrH) + lane);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = warp * R + i0 + i;
          if (r < rows_valid) {                    // (warp-uniform)
            const size_t gr = (size_t)(row0 + r);
            const float4 x = *reinterpret_cast<const float4*>(T + stg_off(r, 4 * lane));
            const float mean = warp_sum((x.x + x.y) + (x.z + x.w)) * (1.0f / kTrH);
            const float4 d = make_float4(x.x - mean, x.y - mean, x.z - mean, x.w - mean);
            const float var = warp_sum(fmaf(d.x, d.x, fmaf(d.y, d.y, fmaf(d.z, d.z, d.w * d.w)))) * (1.0f / kTrH);
            const float rstd = 1.0f / sqrtf(var + 1e-5f);
            const float4 xh = make_float4(d.x * rstd, d.y * rstd, d.z * rstd, d.w * rstd);
            if (a.ln_xhat) reinterpret_cast<float4*>(a.ln_xhat + gr * kTrH)[lane] = xh;
            if (a.ln_rstd && lane == 0) a.ln_rstd[gr] = rstd;
            const float4 y = make_float4(xh.x * gg.x + bb.x, xh.y * gg.y + bb.y, xh.z * gg.z + bb.z, xh.w * gg.w + bb.w);
            if (a.out) reinterpret_cast<float4*>(a.out + gr * kTrH)[lane] = y;
            if (a.res_out)
              reinterpret_cast<float4*>(a.res_out + gr * kTrH)[lane] =
                  make_float4(ee[i].x + y.x, ee[i].y + y.y, ee[i].z + y.z, ee[i].w + y.w);
          }
        }
      }
    } else {
#pragma unroll 1
    for (int i0 = 0; i0 < R; i0 += 4) {
      float4 t1[4], t2[4], mk[4], oo[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = warp * R + i0 + i;
        const bool ok = true;                      // (loads are clamped to the last valid row instead of branched around)
        const size_t gr = (size_t)(row0 + min(r, rows_valid - 1));
        int s1 = 0, s2 = 0;
        if constexpr (kMode == 0) { s1 = __shfl_sync(0xffffffffu, my_i1, i0 + i); s2 = __shfl_sync(0xffffffffu, my_i2, i0 + i); }
        t1[i] = t2[i] = mk[i] = oo[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (has_add1 && ok) t1[i] = __ldg(reinterpret_cast<const float4*>(a.add1 + (size_t)s1 * kTrH) + lane);
        if (has_add2 && ok) t2[i] = __ldg(reinterpret_cast<const float4*>(a.add2 + (size_t)s2 * kTrH) + lane);
        if (has_mask && ok) mk[i] = __ldg(reinterpret_cast<const float4*>(a.mask + gr * kTrH) + lane);
        if (has_acc && ok) oo[i] = *(reinterpret_cast<const float4*>(a.out + gr * kTrH) + lane);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = warp * R + i0 + i;
        {
          float4 v = *reinterpret_cast<const float4*>(T + stg_off(r, 4 * lane));
          if (has_add1) { v.x += t1[i].x; v.y += t1[i].y; v.z += t1[i].z; v.w += t1[i].w; }
          if (has_add2) { v.x += t2[i].x; v.y += t2[i].y; v.z += t2[i].z; v.w += t2[i].w; }
          if (a.relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
          if (has_mask) {
            v.x = mk[i].x > 0.f ? v.x : 0.f; v.y = mk[i].y > 0.f ? v.y : 0.f;
            v.z = mk[i].z > 0.f ? v.z : 0.f; v.w = mk[i].w > 0.f ? v.w : 0.f;
          }
          if (has_acc) { v.x += oo[i].x; v.y += oo[i].y; v.z += oo[i].z; v.w += oo[i].w; }
          if (r < rows_valid) *(reinterpret_cast<float4*>(a.out + (size_t)(row0 + r) * kTrH) + lane) = v;
        }
      }
    }
    }
    named_bar_sync(1 + slot, kNT);             // the buffer is rewritten by the slot's next tile
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(sh->tmem_base, 128 * kGemmSlots);
}

constexpr int kGemmSmem = (1 + kGemmSlots) * kTrChunkBytes + (int)sizeof(GemmSmem) + 1024;

// ---- weight packing (batched): fp32 [128][ld] block -> bf16 hi / lo swizzled K-major images ----
__global__ void tr_pack_kernel(const TrPackJob* __restrict__ jobs) {
  const TrPackJob j = jobs[blockIdx.y];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;      // one thread per (row n, 8-column group)
  if (idx >= kTrH * (kTrH / 8)) return;
  const int n = idx / (kTrH / 8), c8 = (idx % (kTrH / 8)) * 8;
  float v[8];
#pragma unroll
  for (int t = 0; t < 8; ++t)
    v[t] = j.transpose ? j.w[(size_t)(c8 + t) * j.ld + j.col0 + n] : j.w[(size_t)n * j.ld + j.col0 + c8 + t];
  uint4 hi, lo;
  split_bf16x2(v[0], v[1], hi.x, lo.x);
  split_bf16x2(v[2], v[3], hi.y, lo.y);
  split_bf16x2(v[4], v[5], hi.z, lo.z);
  split_bf16x2(v[6], v[7], hi.w, lo.w);
  const uint32_t o = img_off(n, c8, kKBlock);
  *reinterpret_cast<uint4*>((char*)j.chunk + o) = hi;
  *reinterpret_cast<uint4*>((char*)j.chunk + kImage + o) = lo;
}

// ================================================================================================
// wgrad: D[m, n] = sum_r A[r, m] B[r, n].  The row tiles land in shared memory exactly like a K-major operand of
// tr_gemm (128-byte rows of 64 features, SWIZZLE_128B); the MMA reads them through MN-major descriptors, i.e. with the
// feature index as the M / N coordinate and the row index as K: leading byte offset = the distance between the two
// 64-feature planes, stride byte offset = 8 rows.
// ================================================================================================
constexpr int kWgRows = 64;                    // rows per tile (K extent of one pipeline stage)
constexpr int kWgPlane = kWgRows * 128;        // one 64-feature plane of an image
constexpr int kWgImage = 2 * kWgPlane;         // 16 KB
constexpr int kWgStage = 4 * kWgImage;         // A hi, A lo, B hi, B lo
constexpr int kWgPartialRows = kTrH + 1;       // 128 rows of dW + one row of column sums

__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t saddr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo_bytes >> 4) << 16;     // between 64-element groups along M / N
  d |= (uint64_t)(1024 >> 4) << 32;          // between 8-row groups along K
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
  return d;
}

struct WgradSmem {
  uint64_t bar_free[2], bar_done;
  uint32_t tmem_base;
  uint32_t pad;
  float colsum[kNT / 32][kTrH];
};

__global__ void __launch_bounds__(kNT, 1)
tr_wgrad_kernel(const float* __restrict__ Ag, const float* __restrict__ Bg, int M_max, const int* __restrict__ M_dev,
                int want_colsum, float* __restrict__ partial) {
  extern __shared__ char smem_raw[];
  char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  WgradSmem* sh = reinterpret_cast<WgradSmem*>(base + 2 * kWgStage);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, q = warp & 3, hf = warp >> 2, row = q * 32 + lane;
  const uint32_t bar_done = smem_u32(&sh->bar_done);
  if (tid == 0) {
    mbar_init(smem_u32(&sh->bar_free[0]), 1);
    mbar_init(smem_u32(&sh->bar_free[1]), 1);
    mbar_init(bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&sh->tmem_base), 128);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_base;
  const int M = M_dev ? min(*M_dev, M_max) : M_max;
  const int n_tiles = (M + kWgRows - 1) / kWgRows;
  constexpr int R = kWgRows / (kNT / 32);      // 8 rows of A and of B per warp per tile
  float cs[4] = {0.f, 0.f, 0.f, 0.f};
  int it = 0;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
    const int s = it & 1, row0 = tile * kWgRows, rows_valid = min(kWgRows, M - row0);
    char* St = base + s * kWgStage;
    if (it >= 2) mbar_wait(smem_u32(&sh->bar_free[s]), (uint32_t)((it >> 1) - 1) & 1u);   // MMAs that read the stage are done
    float4 xa[R], xb[R];
#pragma unroll
    for (int i = 0; i < R; ++i) {
      const int r = warp * R + i;
      xa[i] = xb[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < rows_valid) {
        xa[i] = __ldg(reinterpret_cast<const float4*>(Ag + (size_t)(row0 + r) * kTrH) + lane);
        xb[i] = __ldg(reinterpret_cast<const float4*>(Bg + (size_t)(row0 + r) * kTrH) + lane);
      }
    }
#pragma unroll
    for (int i = 0; i < R; ++i) {
      const int r = warp * R + i;
      cs[0] += xa[i].x; cs[1] += xa[i].y; cs[2] += xa[i].z; cs[3] += xa[i].w;
      const uint32_t o = img_off(r, 4 * lane, kWgPlane);
      uint2 hi, lo;
      split_bf16x2(xa[i].x, xa[i].y, hi.x, lo.x);
      split_bf16x2(xa[i].z, xa[i].w, hi.y, lo.y);
      *reinterpret_cast<uint2*>(St + o) = hi;
      *reinterpret_cast<uint2*>(St + kWgImage + o) = lo;
      split_bf16x2(xb[i].x, xb[i].y, hi.x, lo.x);
      split_bf16x2(xb[i].z, xb[i].w, hi.y, lo.y);
      *reinterpret_cast<uint2*>(St + 2 * kWgImage + o) = hi;
      *reinterpret_cast<uint2*>(St + 3 * kWgImage + o) = lo;
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t st = smem_u32(St);
#pragma unroll
      for (int ks = 0; ks < kWgRows / 16; ++ks) {
        const uint32_t ko = ks * 2048;           // 16 rows along K
        const uint64_t ah = umma_desc_mn_sw128(st + ko, kWgPlane), al = umma_desc_mn_sw128(st + kWgImage + ko, kWgPlane);
        const uint64_t bh = umma_desc_mn_sw128(st + 2 * kWgImage + ko, kWgPlane);
        const uint64_t bl = umma_desc_mn_sw128(st + 3 * kWgImage + ko, kWgPlane);
        umma_bf16_ss(tmem, ah, bh, kIdescMN, (it | ks) ? 1u : 0u);
        umma_bf16_ss(tmem, al, bh, kIdescMN, 1u);
        umma_bf16_ss(tmem, ah, bl, kIdescMN, 1u);
      }
      umma_commit(smem_u32(&sh->bar_free[s]));
    }
    __syncwarp();
  }
  if (tid == 0 && it > 0) umma_commit(bar_done);
  __syncwarp();
  float* mine = partial + (size_t)blockIdx.x * kWgPartialRows * kTrH;
  if (it > 0) {
    mbar_wait(bar_done, 0);
    tc_fence_after();
    char* T = base;                              // both stages are free: fp32 staging tile for coalesced rows
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int col0 = hf * 64 + h * 32;
      float v[32];
      tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)col0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(T + stg_off(row, col0 + 4 * j)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) sh->colsum[warp][4 * lane + j] = cs[j];
  tc_fence_before();
  __syncthreads();
  for (int r = warp; r < kTrH; r += kNT / 32) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (it > 0) v = *reinterpret_cast<const float4*>(base + stg_off(r, 4 * lane));
    reinterpret_cast<float4*>(mine + (size_t)r * kTrH)[lane] = v;
  }
  if (want_colsum && tid < kTrH) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kNT / 32; ++w) t += sh->colsum[w][tid];
    mine[(size_t)kTrH * kTrH + tid] = t;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

constexpr int kWgradSmem = 2 * kWgStage + (int)sizeof(WgradSmem) + 1024;

// fixed-order sum of the per-CTA partials: dW[m][col0 + n] += sum_cta partial[cta][m][n]; dbias[n] += column sums
__global__ void tr_wgrad_reduce_kernel(const float* __restrict__ partial, int n_cta, float* __restrict__ dW, int ld,
                                       int col0, float* __restrict__ dbias) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= kWgPartialRows * kTrH) return;
  const int r = idx / kTrH, c = idx - r * kTrH;
  if (r == kTrH && dbias == nullptr) return;
  float t = 0.f;
  for (int b = 0; b < n_cta; ++b) t += partial[(size_t)b * kWgPartialRows * kTrH + idx];
  if (r < kTrH) dW[(size_t)r * ld + col0 + c] += t;
  else dbias[c] += t;
}

}  // namespace

int tr_gemm(const TrGemmArgs& a, cudaStream_t stream) {
  const bool ln = a.ln_g != nullptr;
  GM_REQUIRE(a.x != nullptr && a.chunk != nullptr && (a.out != nullptr || ln), "tr_gemm: NULL argument");
  GM_REQUIRE(!ln || (a.ln_b != nullptr && !a.add1 && !a.add2 && !a.mask && !a.relu && !a.accumulate),
             "tr_gemm: the LayerNorm epilogue excludes adds / ReLU / mask / accumulate");
  GM_REQUIRE(!ln || a.res_out == nullptr || a.res_in != nullptr, "tr_gemm: res_out needs res_in");
  if (a.M <= 0) return GM_OK;
  static PerDeviceOnce once;
  const int dev = once.begin();
  if (dev >= 0) {
    GM_CUDA(cudaFuncSetAttribute(tr_gemm_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
    GM_CUDA(cudaFuncSetAttribute(tr_gemm_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
    GM_CUDA(cudaFuncSetAttribute(tr_gemm_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
    GM_CUDA(cudaFuncSetAttribute(tr_gemm_kernel<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
    GM_CUDA(cudaFuncSetAttribute(tr_gemm_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
    once.finish(dev);
  }
  const int grid = min(ceil_div(a.M, kT), kNumSMs);
  const bool adds = a.add1 != nullptr || a.add2 != nullptr, mask = a.mask != nullptr, acc = a.accumulate != 0;
  if (ln) tr_gemm_kernel<true, 1><<<grid, kGemmThreads, kGemmSmem, stream>>>(a);
  else if (!adds && !mask && !acc) tr_gemm_kernel<false, 1><<<grid, kGemmThreads, kGemmSmem, stream>>>(a);
  else if (!adds && mask && !acc) tr_gemm_kernel<false, 2><<<grid, kGemmThreads, kGemmSmem, stream>>>(a);
  else if (!adds && !mask && acc) tr_gemm_kernel<false, 3><<<grid, kGemmThreads, kGemmSmem, stream>>>(a);
  else tr_gemm_kernel<false, 0><<<grid, kGemmThreads, kGemmSmem, stream>>>(a);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("tr_gemm_kernel");
  return GM_OK;
}

int tr_pack(const TrPackJob* jobs_dev, int n_jobs, cudaStream_t stream) {
  if (n_jobs <= 0) return GM_OK;
  tr_pack_kernel<<<dim3(ceil_div(kTrH * (kTrH / 8), 128), n_jobs), 128, 0, stream>>>(jobs_dev);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("tr_pack_kernel");
  return GM_OK;
}

}  // namespace gm

namespace gm {

size_t tr_wgrad_partial_floats() { return (size_t)kNumSMs * kWgPartialRows * kTrH; }

int tr_wgrad(const float* A, const float* B, int M, const int* M_dev, float* dW, int ld, int col0, float* dbias,
             float* partial, cudaStream_t stream) {
  GM_REQUIRE(A != nullptr && B != nullptr && dW != nullptr && partial != nullptr, "tr_wgrad: NULL argument");
  if (M <= 0) return GM_OK;
  static PerDeviceOnce once;
  const int dev = once.begin();
  if (dev >= 0) {
    GM_CUDA(cudaFuncSetAttribute(tr_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgradSmem));
    once.finish(dev);
  }
  const int grid = min(ceil_div(M, kWgRows), kNumSMs);
  tr_wgrad_kernel<<<grid, kNT, kWgradSmem, stream>>>(A, B, M, M_dev, dbias != nullptr ? 1 : 0, partial);
  tr_wgrad_reduce_kernel<<<ceil_div(kWgPartialRows * kTrH, 256), 256, 0, stream>>>(partial, grid, dW, ld, col0, dbias);
  GM_LAUNCHED(2);
  GM_LAUNCH_CHECK("tr_wgrad kernels");
  return GM_OK;
}

}  // namespace gm

// ================================================================================================
// CUDA-core kernels: LayerNorm forward / backward, small first / last layers, loss, Adam
// ================================================================================================
namespace gm {
namespace {

constexpr int kLnWarps = 8;                    // rows per block in the row kernels (warp per row)
constexpr int kLnRowsPerBlock = 64;            // ln_bwd: rows a block walks (its dgamma / dbeta partial)


// torch.nn.LayerNorm(128, eps=1e-5): biased variance, two-pass
__global__ void __launch_bounds__(kLnWarps * 32)
tr_ln_fwd_kernel(const float* __restrict__ z, const float* __restrict__ g, const float* __restrict__ b, int M_max,
                 const int* __restrict__ M_dev, float* __restrict__ xhat, float* __restrict__ rstd_out,
                 float* __restrict__ y_out, const float* __restrict__ res_in, float* __restrict__ res_out) {
  const int M = M_dev ? min(*M_dev, M_max) : M_max;
  const int r = blockIdx.x * kLnWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= M) return;
  const float4 x = reinterpret_cast<const float4*>(z + (size_t)r * kTrH)[lane];
  const float mean = warp_sum((x.x + x.y) + (x.z + x.w)) * (1.0f / kTrH);
  const float4 d = make_float4(x.x - mean, x.y - mean, x.z - mean, x.w - mean);
  const float var = warp_sum(fmaf(d.x, d.x, fmaf(d.y, d.y, fmaf(d.z, d.z, d.w * d.w)))) * (1.0f / kTrH);
  const float rstd = 1.0f / sqrtf(var + 1e-5f);
  const float4 xh = make_float4(d.x * rstd, d.y * rstd, d.z * rstd, d.w * rstd);
  if (xhat) reinterpret_cast<float4*>(xhat + (size_t)r * kTrH)[lane] = xh;
  if (rstd_out && lane == 0) rstd_out[r] = rstd;
  const float4 gg = reinterpret_cast<const float4*>(g)[lane], bb = reinterpret_cast<const float4*>(b)[lane];
  const float4 y = make_float4(xh.x * gg.x + bb.x, xh.y * gg.y + bb.y, xh.z * gg.z + bb.z, xh.w * gg.w + bb.w);
  if (y_out) reinterpret_cast<float4*>(y_out + (size_t)r * kTrH)[lane] = y;
  if (res_out) {
    const float4 e = reinterpret_cast<const float4*>(res_in + (size_t)r * kTrH)[lane];
    reinterpret_cast<float4*>(res_out + (size_t)r * kTrH)[lane] = make_float4(e.x + y.x, e.y + y.y, e.z + y.z, e.w + y.w);
  }
}

// one block walks kLnRowsPerBlock rows (warp per row, 8 at a time) and leaves its dgamma / dbeta partial sums
__global__ void __launch_bounds__(kLnWarps * 32)
tr_ln_bwd_kernel(const float* __restrict__ d1, const float* __restrict__ d2, const int* __restrict__ idx2,
                 const float* __restrict__ xhat, const float* __restrict__ rstd, const float* __restrict__ g, int M_max,
                 const int* __restrict__ M_dev, float* __restrict__ dz, float* __restrict__ partial) {
  __shared__ __align__(16) float s_part[kLnWarps][2][kTrH];
  const int M = M_dev ? min(*M_dev, M_max) : M_max;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float4 gg = reinterpret_cast<const float4*>(g)[lane];
  float4 ag = make_float4(0.f, 0.f, 0.f, 0.f), ab = make_float4(0.f, 0.f, 0.f, 0.f);
  const int base = blockIdx.x * kLnRowsPerBlock;
  for (int i = warp; i < kLnRowsPerBlock; i += kLnWarps) {
    const int r = base + i;
    if (r >= M) break;
    float4 dy = reinterpret_cast<const float4*>(d1 + (size_t)r * kTrH)[lane];
    if (d2) {
      const float4 t = reinterpret_cast<const float4*>(d2 + (size_t)(idx2 ? idx2[r] : r) * kTrH)[lane];
      dy.x += t.x; dy.y += t.y; dy.z += t.z; dy.w += t.w;
    }
    const float4 xh = reinterpret_cast<const float4*>(xhat + (size_t)r * kTrH)[lane];
    ab.x += dy.x; ab.y += dy.y; ab.z += dy.z; ab.w += dy.w;
    ag.x = fmaf(dy.x, xh.x, ag.x); ag.y = fmaf(dy.y, xh.y, ag.y); ag.z = fmaf(dy.z, xh.z, ag.z); ag.w = fmaf(dy.w, xh.w, ag.w);
    const float4 t = make_float4(dy.x * gg.x, dy.y * gg.y, dy.z * gg.z, dy.w * gg.w);
    const float m1 = warp_sum((t.x + t.y) + (t.z + t.w)) * (1.0f / kTrH);
    const float m2 = warp_sum(fmaf(t.x, xh.x, fmaf(t.y, xh.y, fmaf(t.z, xh.z, t.w * xh.w)))) * (1.0f / kTrH);
    const float rs = rstd[r];
    reinterpret_cast<float4*>(dz + (size_t)r * kTrH)[lane] =
        make_float4(rs * (t.x - m1 - xh.x * m2), rs * (t.y - m1 - xh.y * m2), rs * (t.z - m1 - xh.z * m2),
                    rs * (t.w - m1 - xh.w * m2));
  }
  reinterpret_cast<float4*>(s_part[warp][0])[lane] = ag;
  reinterpret_cast<float4*>(s_part[warp][1])[lane] = ab;
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * kTrH; i += blockDim.x) {
    const int which = i / kTrH, c = i - which * kTrH;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kLnWarps; ++w) t += s_part[w][which][c];
    partial[(size_t)blockIdx.x * 2 * kTrH + i] = t;
  }
}

// out[i] += sum_b partial[b][i] in a fixed order (i < width), in two levels: groups of kRedChunk consecutive blocks are
// summed in place into the group's first row (one block of threads per group), then the group rows are summed by one
// thread per column.  n_blocks may come from a device row count.
constexpr int kRedChunk = 64;
__global__ void tr_partial_group_kernel(float* __restrict__ partial, int n_blocks_max, int rows_per_block,
                                        const int* __restrict__ M_dev, int width) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= width) return;
  int nb = n_blocks_max;
  if (M_dev) nb = min(nb, (*M_dev + rows_per_block - 1) / rows_per_block);
  const int b0 = blockIdx.y * kRedChunk, b1 = min(b0 + kRedChunk, nb);
  if (b0 + 1 >= b1) return;
  float t = 0.f;
  for (int b = b0; b < b1; ++b) t += partial[(size_t)b * width + i];
  partial[(size_t)b0 * width + i] = t;
}
__global__ void tr_partial_reduce_kernel(const float* __restrict__ partial, int n_blocks_max, int rows_per_block,
                                         const int* __restrict__ M_dev, int width, float* __restrict__ out0,
                                         float* __restrict__ out1, int split) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= width) return;
  int nb = n_blocks_max;
  if (M_dev) nb = min(nb, (*M_dev + rows_per_block - 1) / rows_per_block);
  float t = 0.f;
  for (int b = 0; b < nb; b += kRedChunk) t += partial[(size_t)b * width + i];
  if (i < split) out0[i] += t;
  else if (out1) out1[i - split] += t;
}
static void launch_partial_reduce(float* partial, int nb, int rows_per_block, const int* M_dev, int width, float* out0,
                                  float* out1, int split, cudaStream_t stream) {
  const int bx = ceil_div(width, 256);
  tr_partial_group_kernel<<<dim3(bx, ceil_div(nb, kRedChunk)), 256, 0, stream>>>(partial, nb, rows_per_block, M_dev, width);
  tr_partial_reduce_kernel<<<bx, 256, 0, stream>>>(partial, nb, rows_per_block, M_dev, width, out0, out1, split);
}

// ---- small layers: x [M, K], w [O][K]; a block handles kSmallRows rows ----
constexpr int kSmallRows = 64;
constexpr int kSmallMaxK = 32, kSmallMaxO = 128;

__global__ void __launch_bounds__(256)
tr_small_fwd_kernel(const float* __restrict__ x, int K, const float* __restrict__ w, const float* __restrict__ b, int O,
                    int M_max, const int* __restrict__ M_dev, int relu, float* __restrict__ h) {
  extern __shared__ float s_small[];
  float* sw = s_small;                       // [O][K]
  float* sx = s_small + O * K;               // [rows][K]
  const int M = M_dev ? min(*M_dev, M_max) : M_max;
  const int row0 = blockIdx.x * kSmallRows, rows = min(kSmallRows, M - row0);
  if (rows <= 0) return;
  for (int i = threadIdx.x; i < O * K; i += blockDim.x) sw[i] = w[i];
  for (int i = threadIdx.x; i < rows * K; i += blockDim.x) sx[i] = x[(size_t)row0 * K + i];
  __syncthreads();
  for (int i = threadIdx.x; i < rows * O; i += blockDim.x) {
    const int r = i / O, o = i - r * O;
    float acc = b ? b[o] : 0.f;
    for (int k = 0; k < K; ++k) acc = fmaf(sx[r * K + k], sw[o * K + k], acc);
    h[(size_t)(row0 + r) * O + o] = relu ? fmaxf(acc, 0.f) : acc;
  }
}

// per block: partial dW [O][K] and db [O] over its rows; dx rows (optional)
__global__ void __launch_bounds__(256)
tr_small_bwd_kernel(const float* __restrict__ x, int K, const float* __restrict__ w, int O, const float* __restrict__ dh,
                    const float* __restrict__ mask, int M_max, const int* __restrict__ M_dev, float* __restrict__ dx,
                    const float* __restrict__ dx_mask, float* __restrict__ partial) {
  extern __shared__ float s_small[];
  float* sx = s_small;                       // [rows][K]
  float* sd = s_small + kSmallRows * K;      // [rows][O]
  float* sw = sd + kSmallRows * O;           // [O][K] (dx only)
  const int M = M_dev ? min(*M_dev, M_max) : M_max;
  const int row0 = blockIdx.x * kSmallRows, rows = max(0, min(kSmallRows, M - row0));
  for (int i = threadIdx.x; i < rows * K; i += blockDim.x) sx[i] = x[(size_t)row0 * K + i];
  for (int i = threadIdx.x; i < rows * O; i += blockDim.x) {
    float d = dh[(size_t)row0 * O + i];
    if (mask && !(mask[(size_t)row0 * O + i] > 0.f)) d = 0.f;
    sd[i] = d;
  }
  if (dx) for (int i = threadIdx.x; i < O * K; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  float* mine = partial + (size_t)blockIdx.x * (O * K + O);
  for (int i = threadIdx.x; i < O * K + O; i += blockDim.x) {
    float acc = 0.f;
    if (i < O * K) {
      const int o = i / K, k = i - o * K;
      for (int r = 0; r < rows; ++r) acc = fmaf(sd[r * O + o], sx[r * K + k], acc);
    } else {
      const int o = i - O * K;
      for (int r = 0; r < rows; ++r) acc += sd[r * O + o];
    }
    mine[i] = acc;
  }
  if (dx) {
    for (int i = threadIdx.x; i < rows * K; i += blockDim.x) {
      const int r = i / K, k = i - r * K;
      float acc = 0.f;
      for (int o = 0; o < O; ++o) acc = fmaf(sd[r * O + o], sw[o * K + k], acc);
      if (dx_mask && !(dx_mask[(size_t)(row0 + r) * K + k] > 0.f)) acc = 0.f;
      dx[(size_t)(row0 + r) * K + k] = acc;
    }
  }
}

// L1Loss(sum) / n over the selected rows: block partials of |pred - tgt| and of the row count, then one finishing block
__global__ void tr_l1_partial_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                     const unsigned char* __restrict__ row_mask, int M, int width,
                                     float* __restrict__ scratch) {
  __shared__ float s_a[8], s_n[8];
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  float a = 0.f, n = 0.f;
  if (r < M && (!row_mask || row_mask[r])) {
    n = 1.f;
    for (int c = 0; c < width; ++c) a += fabsf(pred[(size_t)r * width + c] - tgt[(size_t)r * width + c]);
  }
  a = warp_sum(a);
  n = warp_sum(n);
  if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = a; s_n[threadIdx.x >> 5] = n; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float ta = 0.f, tn = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { ta += s_a[w]; tn += s_n[w]; }
    scratch[2 + 2 * blockIdx.x] = ta;
    scratch[3 + 2 * blockIdx.x] = tn;
  }
}
__global__ void tr_l1_finish_kernel(float* __restrict__ scratch, int n_blocks, float* __restrict__ loss) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float ta = 0.f, tn = 0.f;
  for (int b = 0; b < n_blocks; ++b) { ta += scratch[2 + 2 * b]; tn += scratch[3 + 2 * b]; }
  scratch[0] = ta;
  scratch[1] = tn;
  *loss = tn > 0.f ? ta / tn : 0.f;
}
__global__ void tr_l1_grad_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                  const unsigned char* __restrict__ row_mask, int M, int width,
                                  const float* __restrict__ scratch, float* __restrict__ dpred) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M * width) return;
  const int r = i / width;
  const float n = scratch[1];
  float g = 0.f;
  if (n > 0.f && (!row_mask || row_mask[r])) {
    const float d = pred[i] - tgt[i];
    g = (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) / n;      // torch: sign(0) = 0
  }
  dpred[i] = g;
}

__global__ void tr_adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                               float* __restrict__ v, size_t n, float lr, float beta1, float beta2, float eps,
                               float bc1, float bc2_sqrt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // torch.optim.Adam (single-tensor path): m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2;
  // p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)
  const float gi = g[i];
  const float mi = beta1 * m[i] + (1.f - beta1) * gi;
  const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  p[i] -= (lr / bc1) * mi / (sqrtf(vi) / bc2_sqrt + eps);
}

__global__ void tr_add_rows_kernel(const float* __restrict__ a, const float* __restrict__ b, const int* __restrict__ idx,
                                   int M_max, const int* __restrict__ M_dev, float* __restrict__ out) {
  const int M = M_dev ? min(*M_dev, M_max) : M_max;
  const int r = blockIdx.x * kLnWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= M) return;
  const float4 x = reinterpret_cast<const float4*>(a + (size_t)r * kTrH)[lane];
  const float4 y = reinterpret_cast<const float4*>(b + (size_t)(idx ? idx[r] : r) * kTrH)[lane];
  reinterpret_cast<float4*>(out + (size_t)r * kTrH)[lane] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
}

}  // namespace

int tr_ln_fwd(const float* z, const float* g, const float* b, int M, const int* M_dev, float* xhat, float* rstd,
              float* y_out, const float* res_in, float* res_out, cudaStream_t stream) {
  if (M <= 0) return GM_OK;
  GM_REQUIRE(res_out == nullptr || res_in != nullptr, "tr_ln_fwd: res_out needs res_in");
  tr_ln_fwd_kernel<<<ceil_div(M, kLnWarps), kLnWarps * 32, 0, stream>>>(z, g, b, M, M_dev, xhat, rstd, y_out, res_in, res_out);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("tr_ln_fwd_kernel");
  return GM_OK;
}

size_t tr_ln_partial_floats(int M) { return (size_t)ceil_div(M > 0 ? M : 1, kLnRowsPerBlock) * 2 * kTrH; }

int tr_ln_bwd(const float* d1, const float* d2, const int* idx2, const float* xhat, const float* rstd, const float* g,
              int M, const int* M_dev, float* dz, float* dgamma, float* dbeta, float* partial, cudaStream_t stream) {
  if (M <= 0) return GM_OK;
  const int nb = ceil_div(M, kLnRowsPerBlock);
  tr_ln_bwd_kernel<<<nb, kLnWarps * 32, 0, stream>>>(d1, d2, idx2, xhat, rstd, g, M, M_dev, dz, partial);
  launch_partial_reduce(partial, nb, kLnRowsPerBlock, M_dev, 2 * kTrH, dgamma, dbeta, kTrH, stream);
  GM_LAUNCHED(3);
  GM_LAUNCH_CHECK("tr_ln_bwd kernels");
  return GM_OK;
}

// ---- LayerNorm backward without any caller-provided or allocated scratch (gm_train_layernorm_bwd_static) ----
// The per-block dgamma / dbeta partials live in a file-scope device array: kLnStaticMaxRows rows -> 1 024 blocks x 256
// floats = 1 MB, held by every GPU that loads the library until the process exits.  All calls share it: calls that
// overlap on the GPU (different streams, or two host threads) race on it.  Nothing in it has to start at zero -- every
// block overwrites its whole 256-float slice before the reduction reads it -- so the previous call's contents are harmless.
constexpr int kLnStaticMaxRows = 65536;
constexpr int kLnStaticFloats = (kLnStaticMaxRows / kLnRowsPerBlock) * 2 * kTrH;
static __device__ __align__(256) float g_ln_static_partial[kLnStaticFloats];

// Device address of g_ln_static_partial on the CURRENT device, looked up once per device: in host code the array's
// name is only a host-side stand-in for the symbol, never a pointer a kernel may dereference.
static int ln_static_partial_ptr(float** out) {
  constexpr int kMaxDevices = PerDeviceOnce::kMaxDevices;
  static float* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  GM_CUDA(cudaGetDevice(&dev));
  GM_REQUIRE(dev >= 0 && dev < kMaxDevices, "layernorm_bwd_static: device index %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    float* p = nullptr;
    GM_CUDA(cudaGetSymbolAddress((void**)&p, g_ln_static_partial));
    GM_REQUIRE(p != nullptr, "layernorm_bwd_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return GM_OK;
}

int tr_ln_bwd_static(const float* d1, const float* d2, const int* idx2, const float* xhat, const float* rstd,
                     const float* g, int M, float* dz, float* dgamma, float* dbeta, cudaStream_t stream) {
  GM_REQUIRE(M >= 0 && M <= kLnStaticMaxRows,
             "gm_train_layernorm_bwd_static: n_rows=%d exceeds the static scratch (at most %d rows); use "
             "gm_train_layernorm_bwd with a caller-provided partial buffer", M, kLnStaticMaxRows);
  GM_REQUIRE(d1 != nullptr && xhat != nullptr && rstd != nullptr && g != nullptr && dz != nullptr && dgamma != nullptr &&
                 dbeta != nullptr, "gm_train_layernorm_bwd_static: NULL argument");
  float* partial = nullptr;
  int rc = ln_static_partial_ptr(&partial);
  if (rc) return rc;
  return tr_ln_bwd(d1, d2, idx2, xhat, rstd, g, M, nullptr, dz, dgamma, dbeta, partial, stream);
}

int tr_small_fwd(const float* x, int K, const float* w, const float* b, int O, int M, const int* M_dev, int relu,
                 float* h, cudaStream_t stream) {
  GM_REQUIRE(K >= 1 && K <= kSmallMaxK * 4 && O >= 1 && O <= kSmallMaxO, "tr_small_fwd: K=%d O=%d", K, O);
  if (M <= 0) return GM_OK;
  const size_t smem = sizeof(float) * ((size_t)O * K + (size_t)kSmallRows * K);
  GM_REQUIRE(smem <= 48 * 1024, "tr_small_fwd: K=%d O=%d need %zu bytes of shared memory", K, O, smem);
  tr_small_fwd_kernel<<<ceil_div(M, kSmallRows), 256, smem, stream>>>(x, K, w, b, O, M, M_dev, relu, h);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("tr_small_fwd_kernel");
  return GM_OK;
}

size_t tr_small_partial_floats(int M, int K, int O) { return (size_t)ceil_div(M > 0 ? M : 1, kSmallRows) * (O * K + O); }

int tr_small_bwd(const float* x, int K, const float* w, int O, const float* dh, const float* mask, int M,
                 const int* M_dev, float* dW, float* db, float* dx, const float* dx_mask, float* partial,
                 cudaStream_t stream) {
  if (M <= 0) return GM_OK;
  const size_t smem = sizeof(float) * ((size_t)kSmallRows * K + (size_t)kSmallRows * O + (dx ? (size_t)O * K : 0));
  GM_REQUIRE(smem <= 96 * 1024, "tr_small_bwd: K=%d O=%d need %zu bytes of shared memory", K, O, smem);
  static PerDeviceOnce once;
  const int dev = once.begin();
  if (dev >= 0) {
    GM_CUDA(cudaFuncSetAttribute(tr_small_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    once.finish(dev);
  }
  const int nb = ceil_div(M, kSmallRows), width = O * K + O;
  tr_small_bwd_kernel<<<nb, 256, smem, stream>>>(x, K, w, O, dh, mask, M, M_dev, dx, dx_mask, partial);
  launch_partial_reduce(partial, nb, kSmallRows, M_dev, width, dW, db, O * K, stream);
  GM_LAUNCHED(3);
  GM_LAUNCH_CHECK("tr_small_bwd kernels");
  return GM_OK;
}

int tr_l1_loss(const float* pred, const float* tgt, const unsigned char* row_mask, int M, int width, float* loss,
               float* dpred, float* scratch, cudaStream_t stream) {
  GM_REQUIRE(M > 0 && width > 0, "tr_l1_loss: empty prediction");
  const int nb = ceil_div(M, 256);
  tr_l1_partial_kernel<<<nb, 256, 0, stream>>>(pred, tgt, row_mask, M, width, scratch);
  tr_l1_finish_kernel<<<1, 32, 0, stream>>>(scratch, nb, loss);
  if (dpred) tr_l1_grad_kernel<<<ceil_div((long long)M * width, 256), 256, 0, stream>>>(pred, tgt, row_mask, M, width, scratch, dpred);
  GM_LAUNCHED(dpred ? 3 : 2);
  GM_LAUNCH_CHECK("tr_l1_loss kernels");
  return GM_OK;
}

int tr_adam(float* p, const float* g, float* m, float* v, size_t n, float lr, float beta1, float beta2, float eps,
            int step, cudaStream_t stream) {
  GM_REQUIRE(step >= 1, "tr_adam: step=%d (steps count from 1)", step);
  if (n == 0) return GM_OK;
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float bc2 = 1.f - powf(beta2, (float)step);
  tr_adam_kernel<<<ceil_div((long long)n, 256), 256, 0, stream>>>(p, g, m, v, n, lr, beta1, beta2, eps, bc1, sqrtf(bc2));
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("tr_adam_kernel");
  return GM_OK;
}

int tr_add_rows(const float* a, const float* b, const int* idx, int M, const int* M_dev, float* out, cudaStream_t stream) {
  if (M <= 0) return GM_OK;
  tr_add_rows_kernel<<<ceil_div(M, kLnWarps), kLnWarps * 32, 0, stream>>>(a, b, idx, M, M_dev, out);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("tr_add_rows_kernel");
  return GM_OK;
}

}  // namespace gm
