// Row A4 -- node half of a message-passing step, 16-warp version: one 128-node tile per CTA iteration like
// mp_node_kernel (mp_tc.cu), but 512 threads = (lane quadrant q = warp & 3) x (column QUARTER cq = warp >> 2), i.e.
// thread = one accumulator row x 32 columns instead of x 64.
//
// Replaces NodeModel.forward (deps/torch-graphnet/torch_graphnet/graph_networks.py:63-85), the node residual of
// EncProcDecGNN._process (gnn_manip/models/epd_gnn.py:103) and, after the last step, the decoder (epd_gnn.py:49, 96);
// produces Pr = v' Wr^T + b1 and Ps = v' Ws^T of the next edge MLP.
//
// Why: in a single rollout (2 000 particles = 16 node tiles) the kernel is ONE serial chain per CTA -- five dependent
// GEMMs (MMA-paced, ~2 000 cycles each) with an epilogue after each.  The epilogues
// (TMEM load -> bias / ReLU / LayerNorm -> bf16 hi/lo split -> swizzled operand store, and the row copies in and out) are
// per-thread serial work; four warps per lane quadrant halve it.  Shared-memory plan, weight streaming, operand layout and
// arithmetic are those of mp_node_kernel; LayerNorm partial sums are combined over four column quarters.
#include "mp_tc.cuh"

namespace gm {

namespace {

constexpr int kWorkers16 = 512;              // 16 worker warps: (lane quadrant q = warp & 3) x (column quarter cq = warp >> 2)
constexpr int NT16 = kWorkers16 + 32;         // + warp 16: streams the weights and issues every MMA (see issuer16)
constexpr int kW16 = kWorkers16 / 32;         // 16 warps
constexpr int kR16 = kTile / kW16;            // 8 rows per warp in the row-wise phases
constexpr uint32_t kIdesc16 = umma_idesc_bf16(kTile, kH);

struct Misc16 {
  uint64_t bar_w;
  uint64_t bar_mma;
  uint64_t bar_w2;                            // the second weight chunk of a tile's first layer, loaded up front into A0
  uint32_t tmem_base;
  uint32_t pad;
  alignas(16) float red[2][4][kTile];         // [statistic][column quarter][row]
  alignas(16) float dot[3][4][kTile];         // decoder partial dot products
  alignas(16) float prm[1536];                // b1 b2 b3 gamma beta | decoder b1 b2 w3[3][128] b3[3]
};
constexpr int kQB1 = 0, kQB2 = 128, kQB3 = 256, kQG = 384, kQBeta = 512, kQD1 = 640, kQD2 = 768, kQW3 = 896, kQD3 = 1280;
constexpr int kSmem16 = 3 * kChunkBytes + (int)sizeof(Misc16) + 1024;
static_assert(kSmem16 <= 232448, "mp_node16 shared memory exceeds 227 KB");

struct Ctx16 {
  char *A0, *T, *W;
  uint32_t a0_s, t_s, w_s;
  Misc16* misc;
  uint32_t bar_w, bar_mma, tmem, ph_w, ph_mma, bar_w2, ph_w2;
  int tid, lane, warp, q, cq, row, col0;
};

__device__ __forceinline__ uint32_t a_off16(int row, int col) {
  const int kb = col >> 6, chunk = (col & 63) >> 3;
  return kb * kKBlockBytes + row * 128 + ((chunk ^ (row & 7)) << 4) + ((col & 7) << 1);
}
__device__ __forceinline__ uint32_t s_off16(int row, int col) {
  return (uint32_t)row * 512u + ((((uint32_t)col >> 2) ^ ((uint32_t)row & 7u)) << 4) + (((uint32_t)col & 3u) << 2);
}
__device__ __forceinline__ void store_a8_16(char* a, int row, int col, const float* v) {
  uint4 hi, lo;
  split_bf16x2(v[0], v[1], hi.x, lo.x);
  split_bf16x2(v[2], v[3], hi.y, lo.y);
  split_bf16x2(v[4], v[5], hi.z, lo.z);
  split_bf16x2(v[6], v[7], hi.w, lo.w);
  const uint32_t o = a_off16(row, col);
  *reinterpret_cast<uint4*>(a + o) = hi;
  *reinterpret_cast<uint4*>(a + kImageBytes + o) = lo;
}

// issuer lane: one 64 KB weight chunk -> dst_s, completion on `bar`
__device__ __forceinline__ void load_w16_to(const void* chunk, uint32_t dst_s, uint32_t bar) {
  mbar_expect_tx(bar, kChunkBytes);
#pragma unroll
  for (int i = 0; i < 4; ++i)
    bulk_g2s(dst_s + i * kKBlockBytes, (const char*)chunk + (size_t)i * kKBlockBytes, kKBlockBytes, bar);
}

// all threads: publish the operand in buffer `a_s`, then thread 0 waits for the weights and issues the 24 MMAs
// workers: the operand of the next GEMM is complete (their generic-proxy writes made visible to the async proxy); the
// issuer warp takes it from the barrier on
__device__ __forceinline__ void publish16() {
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
}
__device__ __forceinline__ void sync_workers16() { named_bar_sync(5, kWorkers16); }   // (ids 1..4: the LayerNorm quadrants)

// issuer (lane 0 of warp 16): the 24 MMAs of one GEMM, A images at a_s, weight chunk at w_s
__device__ __forceinline__ void issue_gemm16(const Ctx16& c, uint32_t a_s, uint32_t w_s, bool accumulate) {
  const uint64_t a0 = umma_desc_sw128(a_s), w0 = umma_desc_sw128(w_s);   // base descriptors once, one add per MMA operand
#pragma unroll
  for (int kb = 0; kb < 2; ++kb) {
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      const uint32_t off = kb * kKBlockBytes + ks * 32;
      const uint64_t ah = umma_desc_advance(a0, off), al = umma_desc_advance(a0, kImageBytes + off);
      const uint64_t wh = umma_desc_advance(w0, off), wl = umma_desc_advance(w0, kImageBytes + off);
      umma_bf16_ss(c.tmem, ah, wh, kIdesc16, (accumulate || kb || ks) ? 1u : 0u);
      umma_bf16_ss(c.tmem, ah, wl, kIdesc16, 1u);
      umma_bf16_ss(c.tmem, al, wh, kIdesc16, 1u);
    }
  }
  umma_commit(c.bar_mma);
}
__device__ __forceinline__ void wait_gemm16(Ctx16& c) {
  mbar_wait(c.bar_mma, c.ph_mma);
  c.ph_mma ^= 1u;
  tc_fence_after();
}
__device__ __forceinline__ void load_acc16(const Ctx16& c, float* v) {
  tmem_ld32(c.tmem + ((uint32_t)(c.q * 32) << 16) + (uint32_t)c.col0, v);
  tmem_ld_wait();
}

// 128 fp32 rows (8 per warp, coalesced 512 B each) -> bf16 hi / lo operand images
template <typename RowPtr>
__device__ __forceinline__ void rows_to_a16(const Ctx16& c, char* a, int rows_valid, RowPtr rowptr) {
  float4 x[kR16];
#pragma unroll
  for (int i = 0; i < kR16; ++i) {
    const int r = c.warp * kR16 + i;
    x[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < rows_valid) x[i] = __ldg(reinterpret_cast<const float4*>(rowptr(r)) + c.lane);
  }
#pragma unroll
  for (int i = 0; i < kR16; ++i) {
    const int r = c.warp * kR16 + i;
    uint2 hi, lo;
    split_bf16x2(x[i].x, x[i].y, hi.x, lo.x);
    split_bf16x2(x[i].z, x[i].w, hi.y, lo.y);
    const uint32_t o = a_off16(r, 4 * c.lane);
    *reinterpret_cast<uint2*>(a + o) = hi;
    *reinterpret_cast<uint2*>(a + kImageBytes + o) = lo;
  }
}

// relu(D + bias) -> A0 (this thread's 32 columns of its row)
__device__ __forceinline__ void epi_hidden16(const Ctx16& c, const float* __restrict__ bias_s) {
  float v[32];
  load_acc16(c, v);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 bs = *reinterpret_cast<const float4*>(bias_s + c.col0 + 4 * j);
    v[4 * j] = fmaxf(v[4 * j] + bs.x, 0.f); v[4 * j + 1] = fmaxf(v[4 * j + 1] + bs.y, 0.f);
    v[4 * j + 2] = fmaxf(v[4 * j + 2] + bs.z, 0.f); v[4 * j + 3] = fmaxf(v[4 * j + 3] + bs.w, 0.f);
  }
#pragma unroll
  for (int g8 = 0; g8 < 4; ++g8) store_a8_16(c.A0, c.row, c.col0 + 8 * g8, v + 8 * g8);
}

// T (fp32 staging) -> global rows [row0, row0 + rows_valid) of a [*, 128] matrix
__device__ __forceinline__ void stage_out16(const Ctx16& c, float* __restrict__ out, int row0, int rows_valid) {
  float4 o4[kR16];
#pragma unroll
  for (int i = 0; i < kR16; ++i) o4[i] = *reinterpret_cast<const float4*>(c.T + s_off16(c.warp * kR16 + i, 4 * c.lane));
#pragma unroll
  for (int i = 0; i < kR16; ++i) {
    const int r = c.warp * kR16 + i;
    if (r < rows_valid) reinterpret_cast<float4*>(out + (size_t)(row0 + r) * kH)[c.lane] = o4[i];
  }
}

// this thread's 32 accumulator columns (+ optional bias from global) -> T
__device__ __forceinline__ void acc_to_stage16(const Ctx16& c, const float* __restrict__ bias) {
  float v[32];
  load_acc16(c, v);
  if (bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(bias + c.col0) + j);
      v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j)
    *reinterpret_cast<float4*>(c.T + s_off16(c.row, c.col0 + 4 * j)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}

#ifdef GM_STAMPS
// Debug aid (-DGM_STAMPS): thread 0 of CTA 0 records clock64() along its first tile; gm_debug_stamps() entries 32..55
__device__ unsigned long long g_n16_stamps[24];
#define GM_N16(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) g_n16_stamps[i] = clock64(); } while (0)
#else
#define GM_N16(i) do { } while (0)
#endif

}  // namespace

// chunks: m.wc[0] = W1[:, 0:128] (aggregated edges), m.wc[1] = W1[:, 128:256] (node), m.wc[2], m.wc[3];
// next.wc[3] / next.wc[4] = Wr / Ws of the next edge MLP, or (has_dec) next.wc[0] / next.wc[1] = decoder layers 1 / 2
__global__ void __launch_bounds__(NT16, 1)
mp_node16_kernel(TcMlp m, TcMlp next, int has_dec, int out_dim, float* __restrict__ pred, TcGraph g, TcBuffers b) {
  extern __shared__ char smem_raw[];
  char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  Ctx16 c;
  c.A0 = base; c.T = base + kChunkBytes; c.W = base + 2 * kChunkBytes;
  c.a0_s = smem_u32(c.A0); c.t_s = smem_u32(c.T); c.w_s = smem_u32(c.W);
  c.misc = reinterpret_cast<Misc16*>(base + 3 * kChunkBytes);
  c.tid = threadIdx.x; c.lane = c.tid & 31; c.warp = __shfl_sync(0xffffffffu, c.tid >> 5, 0);   // (provably warp-uniform: see mp_edge16_kernel)
  c.q = c.warp & 3; c.cq = c.warp >> 2;
  c.row = c.q * 32 + c.lane;
  c.col0 = c.cq * 32;
  c.bar_w = smem_u32(&c.misc->bar_w);
  c.bar_mma = smem_u32(&c.misc->bar_mma);
  c.bar_w2 = smem_u32(&c.misc->bar_w2);
  c.ph_w = c.ph_mma = c.ph_w2 = 0;
  if (c.tid == 0) {
    mbar_init(c.bar_w, 1);
    mbar_init(c.bar_mma, 1);
    mbar_init(c.bar_w2, 1);
    fence_mbar_init();
  }
  if (c.warp == 0) {
    tmem_alloc(smem_u32(&c.misc->tmem_base), 128);
    tmem_relinquish();
  }
  float* prm = c.misc->prm;
  for (int i = c.tid; i < kH; i += NT16) {
    prm[kQB1 + i] = m.b1[i]; prm[kQB2 + i] = m.b2[i]; prm[kQB3 + i] = m.b3[i];
    prm[kQG + i] = m.ln_g[i]; prm[kQBeta + i] = m.ln_b[i];
    if (has_dec) { prm[kQD1 + i] = next.b1[i]; prm[kQD2 + i] = next.b2[i]; }
  }
  if (has_dec) {
    for (int i = c.tid; i < 3 * kH; i += NT16) prm[kQW3 + i] = i < out_dim * kH ? next.w3[i] : 0.f;
    if (c.tid < 3) prm[kQD3 + c.tid] = c.tid < out_dim ? next.b3[c.tid] : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  c.tmem = c.misc->tmem_base;

  GM_N16(0);
  const int n_tiles = (g.n_nodes + kTile - 1) / kTile;
  bool waited = false;
  if (c.warp == kW16) {
    // ---------------- issuer warp: streams the weight chunks and issues every MMA, so that no worker warp is held up
    // by the issue of a GEMM (the workers copy rows out while the GEMMs of the tail run).  GEMMs of a tile in order:
    // 0 v W1b^T, 1 agg W1a^T (accumulating), 2 layer 2, 3 layer 3, 4 Wr | decoder 1, 5 Ws | decoder 2.
    // Both chunks of layer 1 are loaded up front: W1b into W, W1a into A0 (the aggregate operand is built in T once GEMM 0
    // has read the v operand out of it); every other chunk goes into W as soon as the GEMM before it has completed.
    // (wait-then-trigger like the workers: one early griddepcontrol.launch_dependents from this warp would release the
    // next edge kernel before this step's edge kernel has finished writing the rows its prologue reads)
    bool triggered = false;
    const void* x1 = has_dec ? next.wc[0] : next.wc[3];
    const void* x2 = has_dec ? next.wc[1] : next.wc[4];
    uint32_t ph_w = 0, ph_w2 = 0, ph_mma = 0;
    if (c.lane == 0 && (int)blockIdx.x < n_tiles) {
      load_w16_to(m.wc[1], c.w_s, c.bar_w);
      load_w16_to(m.wc[0], c.a0_s, c.bar_w2);
    }
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
#pragma unroll 1
      for (int gi = 0; gi < 6; ++gi) {
        __syncthreads();                           // the workers' publish16(): operand gi is complete
        if (c.lane == 0) {
          const bool alt = gi == 1;
          mbar_wait(alt ? c.bar_w2 : c.bar_w, alt ? ph_w2 : ph_w);
          if (alt) ph_w2 ^= 1u;
          else ph_w ^= 1u;
          tc_fence_after();
          issue_gemm16(c, gi < 2 ? c.t_s : c.a0_s, alt ? c.a0_s : c.w_s, gi == 1);
          mbar_wait(c.bar_mma, ph_mma);            // GEMM gi complete: its weight buffer is free
          ph_mma ^= 1u;
          if (!triggered) {
            pdl_wait();
            pdl_launch_dependents();
            triggered = true;
          }
          const void* nxt = gi == 0 ? m.wc[2] : gi == 2 ? m.wc[3] : gi == 3 ? x1 : gi == 4 ? x2 : nullptr;
          if (nxt != nullptr) load_w16_to(nxt, c.w_s, c.bar_w);
          if (gi == 5 && tile + (int)gridDim.x < n_tiles) {
            load_w16_to(m.wc[1], c.w_s, c.bar_w);
            load_w16_to(m.wc[0], c.a0_s, c.bar_w2);
          }
        }
        __syncwarp();
      }
    }
    if (!triggered) {
      pdl_wait();
      pdl_launch_dependents();
    }
  } else {
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int node0 = tile * kTile;
    const int rows_valid = min(kTile, g.n_nodes - node0);
    // ---- layer 1, node part first (needs nothing from this step's edge kernel): v rows -> T as a bf16 operand ----
    {
      const float* v_in = b.v;
      rows_to_a16(c, c.T, rows_valid, [=](int r) { return v_in + (size_t)(node0 + r) * kH; });
    }
    GM_N16(1);
    publish16();
    GM_N16(2);
    const int rp = __ldg(g.recv_ptr + min(node0 + c.warp * kR16 + min(c.lane, kR16), g.n_nodes));
    wait_gemm16(c);
    GM_N16(3);
    if (!waited) {
      pdl_wait();                                  // aggpart / carry of this step's edge kernel are complete from here on
      pdl_launch_dependents();
      waited = true;
    }
    GM_N16(4);
    // ---- aggregate per node (first-tile partial + carries of the following 64-row blocks) -> T ----
    // All eight rows of the warp with every load issued before the first use: the partial row and the first carry (a
    // receiver's segment rarely spans more than two 64-row blocks) are independent predicated loads, not a data-dependent
    // loop per row; summation order as before (partial, then the carries in block order).
    {
      float4 x[kR16], c1[kR16];
      int t0[kR16], t1[kR16];
#pragma unroll
      for (int i = 0; i < kR16; ++i) {
        const int r = c.warp * kR16 + i;
        const int sb = __shfl_sync(0xffffffffu, rp, i);
        int se = __shfl_sync(0xffffffffu, rp, i + 1);
        if (r >= rows_valid) se = sb;
        const bool any = se > sb;
        t0[i] = sb / kAggRows;
        t1[i] = any ? (se - 1) / kAggRows : t0[i] - 1;      // t1 - t0 = number of carries; -1: no edges at all
        x[i] = c1[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (any) x[i] = __ldg(reinterpret_cast<const float4*>(b.aggpart + (size_t)(node0 + r) * kH) + c.lane);
        if (t1[i] >= t0[i] + 1) c1[i] = __ldg(reinterpret_cast<const float4*>(b.carry + (size_t)(t0[i] + 1) * kH) + c.lane);
      }
#pragma unroll
      for (int i = 0; i < kR16; ++i) {
        if (t1[i] >= t0[i] + 1) { x[i].x += c1[i].x; x[i].y += c1[i].y; x[i].z += c1[i].z; x[i].w += c1[i].w; }
        for (int t = t0[i] + 2; t <= t1[i]; ++t) {
          const float4 y = __ldg(reinterpret_cast<const float4*>(b.carry + (size_t)t * kH) + c.lane);
          x[i].x += y.x; x[i].y += y.y; x[i].z += y.z; x[i].w += y.w;
        }
        uint2 hi, lo;
        split_bf16x2(x[i].x, x[i].y, hi.x, lo.x);
        split_bf16x2(x[i].z, x[i].w, hi.y, lo.y);
        const uint32_t o = a_off16(c.warp * kR16 + i, 4 * c.lane);
        *reinterpret_cast<uint2*>(c.T + o) = hi;
        *reinterpret_cast<uint2*>(c.T + kImageBytes + o) = lo;
      }
    }
    GM_N16(5);
    publish16();   // D += agg W1a^T (weights in A0)
    GM_N16(6);
    wait_gemm16(c);
    GM_N16(7);
    {                                              // T is free: the fp32 residual rows travel into it meanwhile
      const float* v_in = b.v;
#pragma unroll
      for (int i = 0; i < kR16; ++i) {
        const int r = c.warp * kR16 + i;
        const bool ok = r < rows_valid;
        cp_async16(c.t_s + s_off16(r, 4 * c.lane), ok ? v_in + (size_t)(node0 + r) * kH + 4 * c.lane : v_in, ok ? 16u : 0u);
      }
      cp_async_commit();
    }
    epi_hidden16(c, prm + kQB1);
    GM_N16(8);
    publish16();               // layer 2
    GM_N16(9);
    wait_gemm16(c);
    GM_N16(10);
    epi_hidden16(c, prm + kQB2);
    cp_async_wait_all();
    GM_N16(11);
    publish16();               // layer 3 (its barrier also publishes the residual rows in T)
    GM_N16(12);
    wait_gemm16(c);
    GM_N16(13);
    // ---- LayerNorm + residual: v' -> T (fp32) and A0 (operand) ----
    {
      float x[32];
      load_acc16(c, x);
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 t = *reinterpret_cast<const float4*>(prm + kQB3 + c.col0 + 4 * j);
        x[4 * j] += t.x; x[4 * j + 1] += t.y; x[4 * j + 2] += t.z; x[4 * j + 3] += t.w;
        s += (x[4 * j] + x[4 * j + 1]) + (x[4 * j + 2] + x[4 * j + 3]);
      }
      c.misc->red[0][c.cq][c.row] = s;
      named_bar_sync(1 + c.q, 128);                // the four warps that share this lane quadrant
      const float* r0 = &c.misc->red[0][0][c.row];
      const float mean = ((r0[0] + r0[kTile]) + (r0[2 * kTile] + r0[3 * kTile])) * (1.0f / kH);
      float ss = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) { x[j] -= mean; ss = fmaf(x[j], x[j], ss); }
      c.misc->red[1][c.cq][c.row] = ss;
      named_bar_sync(1 + c.q, 128);
      const float* r1 = &c.misc->red[1][0][c.row];
      const float var = ((r1[0] + r1[kTile]) + (r1[2 * kTile] + r1[3 * kTile])) * (1.0f / kH);
      const float rstd = 1.0f / sqrtf(var + 1e-5f);
      const bool live = c.row < rows_valid;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 g4 = *reinterpret_cast<const float4*>(prm + kQG + c.col0 + 4 * j);
        const float4 b4 = *reinterpret_cast<const float4*>(prm + kQBeta + c.col0 + 4 * j);
        float4* sp = reinterpret_cast<float4*>(c.T + s_off16(c.row, c.col0 + 4 * j));
        const float4 o = *sp;                      // v rows (residual, epd_gnn.py:103)
        x[4 * j] = live ? x[4 * j] * rstd * g4.x + b4.x + o.x : 0.f;
        x[4 * j + 1] = live ? x[4 * j + 1] * rstd * g4.y + b4.y + o.y : 0.f;
        x[4 * j + 2] = live ? x[4 * j + 2] * rstd * g4.z + b4.z + o.z : 0.f;
        x[4 * j + 3] = live ? x[4 * j + 3] * rstd * g4.w + b4.w + o.w : 0.f;
        *sp = make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]);
      }
#pragma unroll
      for (int g8 = 0; g8 < 4; ++g8) store_a8_16(c.A0, c.row, c.col0 + 8 * g8, x + 8 * g8);
    }
    GM_N16(14);
    publish16();               // v' Wr^T   |   decoder layer 1
    GM_N16(15);
    stage_out16(c, b.v, node0, rows_valid);        // v' rows leave while the GEMM runs
    GM_N16(16);
    wait_gemm16(c);
    GM_N16(17);
    if (!has_dec) {
      sync_workers16();                            // every warp is done reading v' out of T
      acc_to_stage16(c, next.b1);                  // Pr carries b1 of the next edge MLP (mp_edge.cu adds no bias)
      GM_N16(18);
      publish16();             // v' Ws^T (the barrier inside publishes T)
      GM_N16(19);
      stage_out16(c, b.pr, node0, rows_valid);
      wait_gemm16(c);
      GM_N16(20);
      sync_workers16();
      acc_to_stage16(c, nullptr);
      tc_fence_before();
      sync_workers16();
      GM_N16(21);
      stage_out16(c, b.ps, node0, rows_valid);
      GM_N16(22);
    } else {
      // decoder: Linear+ReLU, Linear+ReLU on the tensor cores, Linear(128 -> out_dim) on CUDA cores (epd_gnn.py:49, 96)
      epi_hidden16(c, prm + kQD1);
      publish16();
      wait_gemm16(c);
      float acc[3] = {0.f, 0.f, 0.f};
      {
        float v[32];
        load_acc16(c, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float x = fmaxf(v[j] + prm[kQD2 + c.col0 + j], 0.f);
#pragma unroll
          for (int o = 0; o < 3; ++o) acc[o] = fmaf(x, prm[kQW3 + o * kH + c.col0 + j], acc[o]);
        }
      }
#pragma unroll
      for (int o = 0; o < 3; ++o) c.misc->dot[o][c.cq][c.row] = acc[o];
      tc_fence_before();
      sync_workers16();
      if (c.cq == 0 && c.row < rows_valid) {
        for (int o = 0; o < out_dim; ++o) {
          const float* d = &c.misc->dot[o][0][c.row];
          pred[(size_t)(node0 + c.row) * out_dim + o] = ((d[0] + d[kTile]) + (d[2 * kTile] + d[3 * kTile])) + prm[kQD3 + o];
        }
      }
    }
    tc_fence_before();
    sync_workers16();                              // T, dot[] and D are reused by the next tile
  }
  if (!waited) {
    pdl_wait();
    pdl_launch_dependents();
  }
  }
  tc_fence_before();
  __syncthreads();
  if (c.warp == 0) tmem_dealloc(c.tmem, 128);
}

int tc_debug_stamps_node16(unsigned long long* out_host, int n) {
#ifdef GM_STAMPS
  if (n >= 56) GM_CUDA(cudaMemcpyFromSymbol(out_host + 32, g_n16_stamps, sizeof(unsigned long long) * 24));
#else
  (void)out_host; (void)n;
#endif
  return GM_OK;
}

int tc_configure_node16() {
  static PerDeviceOnce once;
  const int dev = once.begin();
  if (dev == -1) return GM_OK;
  GM_REQUIRE(dev >= 0, "tc_configure_node16: cannot query the current device");
  GM_CUDA(cudaFuncSetAttribute(mp_node16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem16));
  once.finish(dev);
  return GM_OK;
}

int tc_mp_node16(const TcMlp& m, const TcMlp* next_edge, const TcMlp* dec, int out_dim, float* pred, const TcGraph& g,
                 const TcBuffers& b, cudaStream_t stream) {
  GM_REQUIRE((next_edge != nullptr) != (dec != nullptr), "tc_mp_node16: exactly one of next_edge / dec");
  GM_REQUIRE(dec == nullptr || (out_dim >= 1 && out_dim <= 3), "tc_mp_node16: out_dim=%d outside [1,3]", out_dim);
  int rc = tc_configure_node16();
  if (rc) return rc;
  const int grid = min(ceil_div(g.n_nodes, kTile), kNumSMs);
  GM_CUDA(launch_pdl(mp_node16_kernel, grid, NT16, (size_t)kSmem16, stream, 1, m, dec ? *dec : *next_edge, dec ? 1 : 0,
                     out_dim, pred, g, b));
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("mp_node16_kernel");
  return GM_OK;
}

GM_TRACE_DUMP(trace_dump_node16)

}  // namespace gm
