// Row A3 (+ the scatter_add of A4) -- edge half of one message-passing step (and the edge encoder) on CTA pairs.
//
// Replaces EdgeModel.forward (deps/torch-graphnet/torch_graphnet/graph_networks.py:29-50), the residual of
// EncProcDecGNN._process (gnn_manip/models/epd_gnn.py:104) and scatter_sum (utils/gn_utils.py:60-64).
//
// Persistent CTA pairs (clusters of 2, one CTA per SM).  mp_edge_kernel: 384 threads per CTA
//   warps 0-3   compute warpgroup of tile slot 0   } thread = one edge row of a 128-edge tile: operand split,
//   warps 4-7   compute warpgroup of tile slot 1   } bias/ReLU/LayerNorm epilogues, segmented receiver sum
//   warps 8-11  hand their registers to the compute warpgroups (setmaxnreg 40 / 232); lane 0 of warp 8 in the
//               leader CTA issues every MMA of both slots, so no compute warp ever blocks on the tensor pipe
// (enc_edge_pair_kernel: 256 threads, lane 0 of each slot's first warp issues.)
// Each CTA has two 128-edge tiles in flight (one per warpgroup).  A layer of a slot runs as ONE M = 256 GEMM over
// the pair (tcgen05.mma.cta_group::2): 128 rows of A from each CTA's tensor memory, the weight rows split 64 + 64
// between the two CTAs' shared memories.  Halving the weights per CTA is what lets all three layers (bf16 hi + lo,
// 96 KB per CTA) stay resident for the whole kernel: no weight streaming, no weight traffic in L2 after the prologue.
// The issuer waits for each K half of an operand (remote mbarrier arrives from both CTAs), issues its 12 MMAs and
// commits a finished layer to the d_ready barrier of both CTAs (multicast); the two slots' GEMMs interleave on the
// tensor pipe in readiness order while the other slot's warps run their epilogue.
//
// Memory plan
//   TMEM (512 columns): per slot two 128-column regions; a layer reads its A operand (bf16 hi / lo, lane = edge
//     row, two K elements per 32-bit column) from one and accumulates fp32 into the other.  Epilogues write the
//     next A operand with tcgen05.st in place over the accumulator block they have just read -- no swizzled
//     shared-memory operand, no st.shared bandwidth -- and the MMAs read A from TMEM and only weights from smem.
//   SMEM: 96 KB of weights; one 64 KB fp32 staging tile per slot, stored as four 32-column boxes of 128 x 128 B
//     with the 16-byte chunk index XOR (row & 7): conflict-free for row-owner 128-bit access and for column-owner
//     32-bit access.  Each warp stages exactly the 32 rows its own threads own (cp.async, one coalesced 512 B row
//     per instruction), so those hand-offs are warp-local: the Ps[send] gather lands while GEMM 1 runs, the e rows
//     of the slot's NEXT tile land while layers 2-3 run and are converted to the next operand before LayerNorm
//     overwrites the rows.  LayerNorm rows leave as vector reductions into e (red.global.add.v4.f32: e' = e +
//     LN(...), the residual is added in L2 and the fp32 residual never enters the SM a second time).
//   Segmented receiver sum: column-owner sweep over the staged rows (running sums restarted by uniform bit masks,
//     written back in place at segment ends), then one coalesced 512 B store per segment by the warp owning the row.
//   bf16x3: x.w ~= xh.wh + xl.wh + xh.wl, accumulated in fp32 in TMEM.
#include "mp_tc.cuh"

namespace gm {

namespace {

constexpr int kNTE = 256;                    // enc_edge: two compute warpgroups; one elected thread per slot issues the MMAs
constexpr int kNTE3 = 384;                   // mp_edge: two compute warpgroups + the MMA-issuer warpgroup
constexpr int kPieceBytes = 16384;           // one K block (64 columns) of one 128-row bf16 image (self test)
constexpr int kWBlock = 8192;                // this CTA's 64 weight rows of one K block of one image
constexpr int kWLayer = 4 * kWBlock;         // {Wh.k0, Wh.k1, Wl.k0, Wl.k1}
constexpr int kWBytes = 3 * kWLayer;         // all three layers stay resident: 96 KB per CTA of the pair
constexpr int kBoxBytes = kTile * 128;       // one 32-column box of the staging tile
constexpr int kStageBytes = 4 * kBoxBytes;   // 64 KB
constexpr uint32_t kIdescE = umma_idesc_bf16(kTile, kH);          // M = 128 (self test, one CTA)
constexpr uint32_t kIdescPair = umma_idesc_bf16(2 * kTile, kH);   // M = 256 across the CTA pair

struct EdgeShared {
  uint64_t w_full;                           // this CTA's 96 KB of weights landed
  uint64_t a_ready[2];                       // leader CTA: operand of the next GEMM is in TMEM (4 + 4 warp arrivals)
  uint64_t d_ready[2];                       // accumulator complete (multicast tcgen05.commit)
  uint32_t tmem_base;
  uint32_t pad;
  uint64_t a_half[2][2];                     // mp_edge_kernel: K half h of slot s's next operand is in TMEM (4 + 4 arrivals)
  uint32_t segmask[2][4];                    // bit l of word q: row 32q+l starts a new receiver segment
  alignas(16) float prm[4 * kH];             // b2 | b3 | LayerNorm gamma | LayerNorm beta
};
constexpr int kEdgeSmemBytes = kWBytes + 2 * kStageBytes + (int)sizeof(EdgeShared);
static_assert(kEdgeSmemBytes <= 232448, "mp_edge shared memory exceeds 227 KB");

// Staging tile addressing.  Element (row, col): box col >> 5, then row * 128 + ((chunk ^ (row & 7)) << 4) + (col & 3) * 4
// with chunk = (col >> 2) & 7.
//   row owner (thread = row): float4 j of the row sits at  box(j >> 3) + rowoff[j & 7]
//   stager    (warp per row): lane l moves float4 l of the row
__device__ __forceinline__ uint32_t stage_chunk_off(int row, int j) {
  return (uint32_t)(j >> 3) * kBoxBytes + (uint32_t)row * 128u + ((((uint32_t)j & 7u) ^ ((uint32_t)row & 7u)) << 4);
}

// Warp-per-row staging of this warp's 32 rows [r0, r0 + 32) (r0 % 8 == 0), one coalesced 512 B row per cp.async
// instruction.  db[ii]: destination of this lane's float4 in row r0 + ii (ii < 8); row r0 + i sits at db[i & 7] +
// (i & ~7) * 128.  n_valid rows are copied, the rest zero-filled.
__device__ __forceinline__ void stage_dst_init(uint32_t S_s, int r0, int lane, uint32_t* db) {
#pragma unroll
  for (int ii = 0; ii < 8; ++ii) db[ii] = S_s + stage_chunk_off(r0 + ii, lane);
}
// contiguous source rows (src0 = this lane's float4 of row r0)
__device__ __forceinline__ void stage_rows32_contig(const uint32_t* db, const float* src0, int n_valid) {
  if (n_valid >= 32) {
#pragma unroll
    for (int i = 0; i < 32; ++i) cp_async16(db[i & 7] + (i & ~7) * 128, src0 + i * kH, 16u);
  } else {
#pragma unroll 8
    for (int i = 0; i < 32; ++i) cp_async16(db[i & 7] + (i & ~7) * 128, i < n_valid ? src0 + i * kH : src0, i < n_valid ? 16u : 0u);
  }
  cp_async_commit();
}
// gathered source rows: lane i of the warp holds the source row index of row r0 + i (< 0: zero fill)
__device__ __forceinline__ void stage_rows32_gather(const uint32_t* db, const float* table_lane, int my_idx) {
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const int sid = __shfl_sync(0xffffffffu, my_idx, i);
    cp_async16(db[i & 7] + (i & ~7) * 128, table_lane + (size_t)(sid < 0 ? 0 : sid) * kH, sid < 0 ? 0u : 16u);
  }
  cp_async_commit();
}

__device__ __forceinline__ void red_add_v4(float* gdst, const float4& v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(gdst), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// ---- CTA-pair (cluster of 2) helpers ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t dst_smem, uint32_t ncols) {  // one warp of EACH CTA
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D (+)= A x B^T over the CTA pair: M = 256 (128 rows of A from each CTA's TMEM), B rows split 64 + 64 between the
// two CTAs' shared memories at the same offset, D rows land in each CTA's own TMEM
__device__ __forceinline__ void umma_bf16_ts_pair(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      :
      : "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {   // arrives on `bar` in both CTAs
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}
// one whole layer GEMM of a slot (24 MMAs): w_layer_s = this CTA's {Wh.k0, Wh.k1, Wl.k0, Wl.k1} blocks of the layer
__device__ __forceinline__ void issue_layer_pair(uint32_t w_layer_s, uint32_t tD, uint32_t tAH, uint32_t tAL) {
  const uint64_t w0 = umma_desc_sw128(w_layer_s);      // base descriptor once, one add per MMA (umma_desc_advance)
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const uint32_t kcol = (uint32_t)(p & 1) * 32u;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks)
      umma_bf16_ts_pair(tD, tAH + kcol + ks * 8, umma_desc_advance(w0, p * kWBlock + ks * 32), kIdescPair, (p | ks) ? 1u : 0u);
    if (p < 2) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks)
        umma_bf16_ts_pair(tD, tAL + kcol + ks * 8, umma_desc_advance(w0, p * kWBlock + ks * 32), kIdescPair, 1u);
    }
  }
}

// 32 lanes x 16 columns: thread i of the warp writes TMEM lane (base_lane + i), columns [col, col + 16)
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem_d] (+)= A[tmem_a] x B[smem]^T, A from tensor memory (lane = row, 8 columns = 16 bf16 along K)
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      :
      : "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// one GEMM piece: p = 0/1 -> Wh K-block p against A hi and A lo; p = 2/3 -> Wl K-block p-2 against A hi
__device__ __forceinline__ void issue_piece(int p, uint32_t w_piece_s, uint32_t tD, uint32_t tAH, uint32_t tAL) {
  const uint32_t kcol = (uint32_t)(p & 1) * 32u;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks)
    umma_bf16_ts(tD, tAH + kcol + ks * 8, umma_desc_sw128(w_piece_s + ks * 32), kIdescE, (p | ks) ? 1u : 0u);
  if (p < 2) {
#pragma unroll
    for (int ks = 0; ks < 4; ++ks)
      umma_bf16_ts(tD, tAL + kcol + ks * 8, umma_desc_sw128(w_piece_s + ks * 32), kIdescE, 1u);
  }
}

// 32 fp32 values -> bf16 hi / lo pairs -> 16 + 16 TMEM columns of the A operand
__device__ __forceinline__ void put_operand32(const float* v, uint32_t tAH, uint32_t tAL, int c32) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) split_bf16x2(v[2 * j], v[2 * j + 1], hi[j], lo[j]);
  tmem_st16(tAH + c32 * 16, hi);
  tmem_st16(tAL + c32 * 16, lo);
}

// this thread's staged fp32 row -> A operand (hi / lo) in TMEM.  S: staging tile, ro[jj]: row-owner chunk offsets
__device__ __forceinline__ void row_to_operand(const char* S, const uint32_t* ro, uint32_t tAH, uint32_t tAL) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float v[32];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 t = *reinterpret_cast<const float4*>(S + c * kBoxBytes + ro[j]);
      v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
    }
    put_operand32(v, tAH, tAL, c);
  }
}

// this warp's rows of the next A operand are in TMEM: tell the leader CTA's issuer (cluster-space barrier address)
__device__ __forceinline__ void publish_operand(uint32_t bar_a_leader, int lane) {
  tmem_st_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive_cluster(bar_a_leader);
}

// ---- in-place operand layout (mp_edge_kernel): a slot owns two 128-column regions X and Y.  A layer reads its A
// operand from one and accumulates into the other; the epilogue then overwrites each 32-column block of the
// accumulator it has just read with the next operand for the same 32 K elements (16 columns bf16 hi + 16 columns
// lo), so the next layer's MMAs over K elements 0..63 can run while the epilogue still works on columns 64..127. ----
__device__ __forceinline__ void put_operand32_inplace(const float* v, uint32_t tR, int c32) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) split_bf16x2(v[2 * j], v[2 * j + 1], hi[j], lo[j]);
  tmem_st16(tR + c32 * 32, hi);
  tmem_st16(tR + c32 * 32 + 16, lo);
}
// the 12 MMAs of one K half (64 K elements) of a layer: A (in-place layout) from region tA, weights {Wh.k, Wl.k}
__device__ __forceinline__ void issue_half_pair(uint32_t w_layer_s, uint32_t tD, uint32_t tA, int half) {
  // base descriptors once, one add per MMA (umma_desc_advance)
  const uint64_t dh = umma_desc_sw128(w_layer_s + half * kWBlock), dl = umma_desc_advance(dh, 2 * kWBlock);
  const uint32_t a0 = tA + 64u * half;         // ks = 4 half + kk: column 32 (ks >> 1) + 8 (ks & 1)
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    const uint32_t a_hi = a0 + 32u * (kk >> 1) + 8u * (kk & 1), a_lo = a_hi + 16u;
    const uint64_t bh = umma_desc_advance(dh, kk * 32), bl = umma_desc_advance(dl, kk * 32);
    umma_bf16_ts_pair(tD, a_hi, bh, kIdescPair, (half | kk) ? 1u : 0u);
    umma_bf16_ts_pair(tD, a_lo, bh, kIdescPair, 1u);
    umma_bf16_ts_pair(tD, a_hi, bl, kIdescPair, 1u);
  }
}

// Phase timestamps (debug aid, -DGM_STAMPS): thread 0 of CTA 0 (slot 0) records clock64() during its third tile,
// the MMA issuer during the matching pair; read back with gm_debug_stamps() (entries 0..31).
#ifdef GM_STAMPS
#ifndef GM_STAMP_TID
#define GM_STAMP_TID 0
#endif
__device__ unsigned long long g_edge_stamps[32];
#define GM_ESTAMP(i) do { if (stamp_on) g_edge_stamps[i] = clock64(); } while (0)
#else
#define GM_ESTAMP(i) do { } while (0)
#endif

constexpr int kOffB2 = 0, kOffB3 = 128, kOffG = 256, kOffBeta = 384;
#define GM_PRM4(ptr) (*reinterpret_cast<const float4*>(ptr))

}  // namespace

__global__ void __launch_bounds__(kNTE3, 1) mp_edge_kernel(TcMlp m, TcGraph g, TcBuffers b) {
  extern __shared__ __align__(1024) char smem_raw[];
  char* base = smem_raw;
  const uint32_t W_s = smem_u32(base);
  char* S_base = base + kWBytes;
  EdgeShared* sh = reinterpret_cast<EdgeShared*>(S_base + 2 * kStageBytes);
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;   // (warp-uniform)
  const uint32_t rank = cluster_ctarank();     // 0: leader (issues the pair's MMAs), 1: peer
  if (tid == 0 && (W_s & 1023u) != 0u) {
    printf("gnn_manip_b200: mp_edge dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  if (tid == 0) {
    mbar_init(smem_u32(&sh->w_full), 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&sh->a_half[s][0]), 8);
      mbar_init(smem_u32(&sh->a_half[s][1]), 8);
      mbar_init(smem_u32(&sh->d_ready[s]), 1);
    }
    fence_mbar_init();
    // this CTA's half (rows 64 rank .. 64 rank + 63) of every K block of every image of the three layers
    mbar_expect_tx(smem_u32(&sh->w_full), kWBytes);
    for (int l = 0; l < 3; ++l)
      for (int p = 0; p < 4; ++p)
        bulk_g2s(W_s + l * kWLayer + p * kWBlock, (const char*)m.wc[l] + (size_t)p * kPieceBytes + rank * kWBlock,
                 kWBlock, smem_u32(&sh->w_full));
  }
  if (warp == 0) tmem_alloc_pair(smem_u32(&sh->tmem_base), 512);
  for (int i = tid; i < kH; i += kNTE3) {
    sh->prm[kOffB2 + i] = m.b2[i];
    sh->prm[kOffB3 + i] = m.b3[i];
    sh->prm[kOffG + i] = m.ln_g[i];
    sh->prm[kOffBeta + i] = m.ln_b[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_base;
                                 // everything above touched constant parameters only

  const int E = min(*g.n_edges_dev, g.max_edges);
  const int n_tiles = (E + kTile - 1) / kTile;
  const int grid = (int)gridDim.x;
  // both CTAs of a pair run the same number of tile iterations (the leader's count); a CTA without a tile for an
  // iteration still takes part in every hand-off with zero valid rows
  const int b_lead = (int)blockIdx.x - (int)rank;
  const int n_my = b_lead < n_tiles ? (n_tiles - b_lead + grid - 1) / grid : 0;
  if (warp >= 8) {
    // ---------------- third warpgroup: gives its registers to the compute warpgroups; its first lane in the
    //                  leader CTA issues every MMA of both slots (nothing else ever blocks on the tensor pipe) ----
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    mbar_wait(smem_u32(&sh->w_full), 0);
    cluster_sync_all();                        // pairs with the compute warpgroups' prologue barrier
    if (warp == 8 && lane == 0 && rank == 0) {
      // items of a slot, in order: (tile, layer, K half); a_half[s][h] is used once per layer -> parity (item / 2) & 1
      int it[2] = {0, 0};
      const int end[2] = {6 * ((n_my + 1) / 2), 6 * (n_my / 2)};
      uint32_t idle = 0;
      while (it[0] < end[0] || it[1] < end[1]) {
        bool progress = false;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          if (it[s] < end[s]) {
            const int half = it[s] & 1, layer = (it[s] >> 1) % 3;
            if (mbar_try_wait(smem_u32(&sh->a_half[s][half]), (uint32_t)(it[s] >> 1) & 1u)) {
              tc_fence_after();
              const uint32_t tXs = tmem + 256u * s, tYs = tXs + 128u;
              issue_half_pair(W_s + layer * kWLayer, layer == 1 ? tXs : tYs, layer == 1 ? tYs : tXs, half);
              if (half) umma_commit_pair(smem_u32(&sh->d_ready[s]));
              ++it[s];
              progress = true;
            }
          }
        }
        if (progress) idle = 0;
        else if (++idle > (1u << 28)) {
          printf("gnn_manip_b200: mp_edge issuer timed out (block %d)\n", blockIdx.x);
          __trap();
        }
      }
    }
    __syncwarp();
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    // ---------------- compute warpgroups: slot s, thread = edge row r ----------------
    const int s = warp >> 2, q = warp & 3, r = tid & (kTile - 1), r0 = q * 32;
    const int KSTEP = 2, K0 = s;             // slot s takes this CTA's tiles s, s + 2, ...
    char* S = S_base + s * kStageBytes;
    const uint32_t S_s = smem_u32(S);
    uint32_t ro[8];                            // row-owner offsets of the 8 chunks of a box
#pragma unroll
    for (int j = 0; j < 8; ++j) ro[j] = stage_chunk_off(r, j);
    // regions of this slot: layer 1 reads X and accumulates into Y, layer 2 reads Y -> X, layer 3 reads X -> Y
    const uint32_t tXs = tmem + 256u * s, tYs = tXs + 128u;                      // MMA addresses (lane 0)
    const uint32_t tX = tXs + ((uint32_t)r0 << 16), tY = tYs + ((uint32_t)r0 << 16);   // this warp's lanes
    const uint32_t bar_a0_local = smem_u32(&sh->a_half[s][0]), bar_a1_local = smem_u32(&sh->a_half[s][1]);
    const uint32_t bar_d = smem_u32(&sh->d_ready[s]);
    const uint32_t bar_a0 = mapa_shared(bar_a0_local, 0), bar_a1 = mapa_shared(bar_a1_local, 0);   // leader CTA's
    const float* prm = sh->prm;
    uint32_t ph_d = 0;
    // leader, one thread per slot: once both CTAs have published the operand, issue the layer's 24 MMAs for the
    // pair and commit to d_ready in both CTAs.  The two slots' issuers interleave on the tensor pipe in readiness order.
    // this warp's part of K half h of the next operand is in TMEM (the issuer warp takes it from there)
    auto publish_half = [&](int half) {
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(half ? bar_a1 : bar_a0);
    };
    auto publish_both = [&]() {
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive_cluster(bar_a0);
        mbar_arrive_cluster(bar_a1);
      }
    };

    int nx_recv = -1, nx_send = -1, nx_prev = -2;
    auto fetch_ids = [&](int k) {
      const int ge = (blockIdx.x + k * grid) * kTile + r;
      nx_recv = nx_send = -1;
      nx_prev = -2;
      if (ge < E) {
        nx_recv = __ldg(g.erecv + ge);
        nx_send = __ldg(g.esend + ge);
        if (ge > 0) nx_prev = __ldg(g.erecv + ge - 1);
      }
    };
    uint32_t db[8];                            // stager destinations of this lane
    stage_dst_init(S_s, r0, lane, db);
    auto stage_e_rows = [&](int k) {           // this warp's 32 rows of tile k
      const int ge0 = (blockIdx.x + k * grid) * kTile + r0;
      stage_rows32_contig(db, b.e + (size_t)(ge0 < E ? ge0 : 0) * kH + 4 * lane, E - ge0);
    };
    int cur_recv = -1, cur_send = -1, cur_prev = -2;
    const float* pr_row = b.pr;
    // e rows of a tile (already landed in this warp's staging rows) -> bf16 hi / lo operand of layer 1 in TMEM
    auto operand_from_stage = [&]() {
      cp_async_wait_all();
      __syncwarp();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float v[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 t = *reinterpret_cast<const float4*>(S + c * kBoxBytes + ro[j]);
          v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
        }
        put_operand32_inplace(v, tX, c);
      }
      tmem_st_wait();
      __syncwarp();                            // every lane is done reading this warp's rows
    };
    // Starts tile k: its layer-1 operand is in TMEM and the staging rows are free.  Segment mask, Pr prefetch,
    // the Ps[send] gather into the staging rows (lands while GEMM 1 runs), operand hand-off, GEMM 1.
    auto begin_tile = [&](int k, bool issue) {
      cur_recv = nx_recv; cur_send = nx_send; cur_prev = nx_prev;
      const int ge = (blockIdx.x + k * grid) * kTile + r;
      {
        const unsigned mk = __ballot_sync(0xffffffffu, cur_recv != cur_prev);
        if (lane == 0) sh->segmask[s][q] = mk;
      }
      pr_row = b.pr + (size_t)(ge < E ? cur_recv : 0) * kH;
#pragma unroll
      for (int i = 0; i < 4; ++i) prefetch_l1(pr_row + 32 * i);
      stage_rows32_gather(db, b.ps + 4 * lane, cur_send);
      if (issue) publish_both();                 // (inside the loop GEMM 1 was already issued under LayerNorm)
    };
    if (K0 < n_my) {
      fetch_ids(K0);
      stage_e_rows(K0);
    }
    // weights of this CTA landed; peer barriers initialised, peer weights landed, TMEM of the pair allocated
    mbar_wait(smem_u32(&sh->w_full), 0);
    cluster_sync_all();
    if (K0 < n_my) {
      operand_from_stage();
      publish_both();                            // layer 1 of the first tile needs e and weights only: its GEMM runs
                                                 // before the dependency wait as well
    }
    // Everything so far read only constants and e, which the kernel two launches back (the previous edge kernel /
    // the edge encoder) finished writing before this grid could start; Pr / Ps come from the node kernel right before.
#ifdef GM_STAMPS
    if (blockIdx.x == 0 && tid == 0) g_edge_stamps[23] = clock64();
#endif
    pdl_wait();
    pdl_launch_dependents();
#ifdef GM_STAMPS
    if (blockIdx.x == 0 && tid == 0) g_edge_stamps[24] = clock64();
#endif
    if (K0 < n_my) begin_tile(K0, false);
    for (int k = K0; k < n_my; k += KSTEP) {
      const int row0 = (blockIdx.x + k * grid) * kTile;
      const int rows_valid = min(kTile, E - row0);
      const bool has_next = k + KSTEP < n_my;
#ifdef GM_STAMPS
      const bool stamp_on = blockIdx.x == 0 && tid == GM_STAMP_TID && k == 4;
#endif
      GM_ESTAMP(0);
      if (has_next) fetch_ids(k + KSTEP);
      // ---- layer 1 epilogue: relu(D + Pr[recv] (b1 folded in) + Ps[send]) ----
      cp_async_wait_all();
      __syncwarp();
      GM_ESTAMP(4);
      mbar_wait(bar_d, ph_d);
      ph_d ^= 1u;
      tc_fence_after();
      GM_ESTAMP(5);
      {
        // accumulator blocks are fetched one ahead (tcgen05.ld is asynchronous; the wait sits after the math of
        // the previous block), Pr comes from L1 (prefetched at the top of the tile)
        float va[32], vb[32];
        float4 pa[8], pb[8];
        auto load_pr = [&](int c, float4* p) {
#pragma unroll
          for (int j = 0; j < 8; ++j) p[j] = __ldg(reinterpret_cast<const float4*>(pr_row + c * 32) + j);
        };
        auto e1_block = [&](int c, float* v, const float4* p) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 s4 = *reinterpret_cast<const float4*>(S + c * kBoxBytes + ro[j]);
            v[4 * j] = fmaxf(v[4 * j] + (p[j].x + s4.x), 0.f);
            v[4 * j + 1] = fmaxf(v[4 * j + 1] + (p[j].y + s4.y), 0.f);
            v[4 * j + 2] = fmaxf(v[4 * j + 2] + (p[j].z + s4.z), 0.f);
            v[4 * j + 3] = fmaxf(v[4 * j + 3] + (p[j].w + s4.w), 0.f);
          }
          put_operand32_inplace(v, tY, c);
        };
        tmem_ld32(tY, va);
        load_pr(0, pa);
        load_pr(1, pb);
        tmem_ld_wait();
        GM_ESTAMP(16);
        tmem_ld32(tY + 32, vb);
        e1_block(0, va, pa);
        GM_ESTAMP(17);
        load_pr(2, pa);
        tmem_ld_wait();
        tmem_ld32(tY + 64, va);
        e1_block(1, vb, pb);
        publish_half(0);                           // K elements 0..63 of the layer-2 operand
        GM_ESTAMP(18);
        load_pr(3, pb);
        tmem_ld_wait();
        tmem_ld32(tY + 96, vb);
        e1_block(2, va, pa);
        GM_ESTAMP(19);
        tmem_ld_wait();
        e1_block(3, vb, pb);
        GM_ESTAMP(20);
      }
      publish_half(1);
      GM_ESTAMP(21);
      // the staging rows are free until the LayerNorm rows go in: the e rows of this slot's next tile land here
      // while layers 2 and 3 run (issued before the issuer blocks on the other warps' hand-off)
      if (has_next) stage_e_rows(k + KSTEP);
      GM_ESTAMP(22);
      GM_ESTAMP(6);
      // ---- layer 2 epilogue: relu(D + b2) ----
      mbar_wait(bar_d, ph_d);
      ph_d ^= 1u;
      tc_fence_after();
      GM_ESTAMP(7);
      {
        float va[32], vb[32];
        auto e2_block = [&](int c, float* v) {
          float4 bs[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) bs[j] = GM_PRM4(prm + kOffB2 + c * 32 + 4 * j);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            v[4 * j] = fmaxf(v[4 * j] + bs[j].x, 0.f);
            v[4 * j + 1] = fmaxf(v[4 * j + 1] + bs[j].y, 0.f);
            v[4 * j + 2] = fmaxf(v[4 * j + 2] + bs[j].z, 0.f);
            v[4 * j + 3] = fmaxf(v[4 * j + 3] + bs[j].w, 0.f);
          }
          put_operand32_inplace(v, tX, c);
        };
        tmem_ld32(tX, va);
        tmem_ld_wait();
        tmem_ld32(tX + 32, vb);
        e2_block(0, va);
        tmem_ld_wait();
        tmem_ld32(tX + 64, va);
        e2_block(1, vb);
        publish_half(0);
        tmem_ld_wait();
        tmem_ld32(tX + 96, vb);
        e2_block(2, va);
        tmem_ld_wait();
        e2_block(3, vb);
      }
      publish_half(1);
      GM_ESTAMP(8);
      // ---- layer 3 epilogue: y = LayerNorm(D + b3) -> staging row ----
      mbar_wait(bar_d, ph_d);
      ph_d ^= 1u;
      tc_fence_after();
      GM_ESTAMP(9);
      if (has_next) operand_from_stage();      // the A region is free (GEMM 3 is complete); D is read right below
      GM_ESTAMP(2);
      {
        // the whole accumulator row lives in registers: exact two-pass LayerNorm, thread-local
        float x[kH];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld32(tY + c * 32, x + c * 32);
        tmem_ld_wait();
        // the accumulator is in registers and the next tile's layer-1 operand is in TMEM: its GEMM 1 runs under the
        // LayerNorm / residual / segmented-sum phases of this tile (E1 of the next tile finds it complete)
        if (has_next) publish_both();              // (the issuer picks it up after its own LayerNorm, below)
        GM_ESTAMP(25);
        float sa = 0.f, sb = 0.f, sc = 0.f, sd = 0.f;
        // parameters are fetched a whole 32-column box at a time (8 independent 128-bit loads in flight) so that
        // their shared-memory latency is paid once per box, not once per 4 columns
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float4 t[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) t[j] = GM_PRM4(prm + kOffB3 + c * 32 + 4 * j);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float* xx = x + c * 32 + 4 * j;
            xx[0] += t[j].x; xx[1] += t[j].y; xx[2] += t[j].z; xx[3] += t[j].w;
            sa += xx[0]; sb += xx[1]; sc += xx[2]; sd += xx[3];
          }
        }
        const float mean = ((sa + sb) + (sc + sd)) * (1.0f / kH);
        GM_ESTAMP(26);
        float qa = 0.f, qb = 0.f, qc = 0.f, qd = 0.f;
#pragma unroll
        for (int j = 0; j < kH / 4; ++j) {
          x[4 * j] -= mean; x[4 * j + 1] -= mean; x[4 * j + 2] -= mean; x[4 * j + 3] -= mean;
          qa = fmaf(x[4 * j], x[4 * j], qa); qb = fmaf(x[4 * j + 1], x[4 * j + 1], qb);
          qc = fmaf(x[4 * j + 2], x[4 * j + 2], qc); qd = fmaf(x[4 * j + 3], x[4 * j + 3], qd);
        }
        const float rstd = 1.0f / sqrtf(((qa + qb) + (qc + qd)) * (1.0f / kH) + 1e-5f);
        GM_ESTAMP(27);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float4 g4[8], b4[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            g4[j] = GM_PRM4(prm + kOffG + c * 32 + 4 * j);
            b4[j] = GM_PRM4(prm + kOffBeta + c * 32 + 4 * j);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float* xx = x + c * 32 + 4 * j;
            *reinterpret_cast<float4*>(S + c * kBoxBytes + ro[j]) =
                make_float4(xx[0] * rstd * g4[j].x + b4[j].x, xx[1] * rstd * g4[j].y + b4[j].y,
                            xx[2] * rstd * g4[j].z + b4[j].z, xx[3] * rstd * g4[j].w + b4[j].w);
          }
        }
      }
      __syncwarp();
      GM_ESTAMP(10);
      // ---- e' = e + y (epd_gnn.py:104): this warp's rows leave as 512 B vector reductions, the add happens in L2 ----
      {
        const int nr = min(32, rows_valid - r0);
        float* e_out = b.e + (size_t)(row0 + r0) * kH + 4 * lane;
        if (nr >= 32) {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            red_add_v4(e_out + i * kH, *reinterpret_cast<const float4*>(S + (db[i & 7] - S_s) + (i & ~7) * 128));
        } else {
#pragma unroll 8
          for (int i = 0; i < 32; ++i)
            if (i < nr) red_add_v4(e_out + i * kH, *reinterpret_cast<const float4*>(S + (db[i & 7] - S_s) + (i & ~7) * 128));
        }
      }
      named_bar_sync(1 + s, kTile);
      GM_ESTAMP(11);
      // ---- receiver-segmented sum of the pre-residual rows (scatter_sum, gn_utils.py:60-64): thread = column r.
      //      Sweep A: running sum along the rows of each 64-row half, restarted at segment starts and written back
      //      in place at segment ends (uniform bit masks, no data-dependent branches, strictly sequential fp32 order).
      //      Sweep B: one store per segment end -> aggpart[receiver] (or carry[] for a segment that began before
      //      this 64-row block). ----
      {
        const uint32_t cbase = (uint32_t)(r >> 5) * kBoxBytes + ((uint32_t)r & 3u) * 4u;
        const uint32_t cchunk = ((uint32_t)r >> 2) & 7u;
        uint32_t co[8];
#pragma unroll
        for (int ii = 0; ii < 8; ++ii) co[ii] = cbase + ((cchunk ^ (uint32_t)ii) << 4);
        const int n0 = max(0, min(kAggRows, rows_valid)), n1 = max(0, min(kAggRows, rows_valid - kAggRows));
        const unsigned long long f0 = ((unsigned long long)sh->segmask[s][1] << 32) | sh->segmask[s][0];
        const unsigned long long f1 = ((unsigned long long)sh->segmask[s][3] << 32) | sh->segmask[s][2];
        const unsigned long long v0 = n0 >= 64 ? ~0ull : ((1ull << n0) - 1ull), v1 = n1 >= 64 ? ~0ull : ((1ull << n1) - 1ull);
        const unsigned long long m0 = f0 & ~1ull & v0, m1 = f1 & ~1ull & v1;              // restart points
        const unsigned long long e0 = ((m0 >> 1) | (n0 > 0 ? 1ull << (n0 - 1) : 0ull)) & v0;   // segment ends
        const unsigned long long e1 = ((m1 >> 1) | (n1 > 0 ? 1ull << (n1 - 1) : 0ull)) & v1;
        float a0 = 0.f, a1 = 0.f;
#pragma unroll 1
        for (int bb = 0; bb < kAggRows / 8; ++bb) {
          float y0[8], y1[8];
          char* p = S + bb * 1024;
#pragma unroll
          for (int ii = 0; ii < 8; ++ii) {
            y0[ii] = *reinterpret_cast<const float*>(p + co[ii] + ii * 128);
            y1[ii] = *reinterpret_cast<const float*>(p + kAggRows * 128 + co[ii] + ii * 128);
          }
          const unsigned sb0 = (unsigned)(m0 >> (8 * bb)) & 0xffu, eb0 = (unsigned)(e0 >> (8 * bb)) & 0xffu;
          const unsigned sb1 = (unsigned)(m1 >> (8 * bb)) & 0xffu, eb1 = (unsigned)(e1 >> (8 * bb)) & 0xffu;
          if ((sb0 | eb0) == 0u) {
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) a0 += y0[ii];
          } else {
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) {
              const float t = a0 + y0[ii];
              a0 = ((sb0 >> ii) & 1u) ? y0[ii] : t;
              if ((eb0 >> ii) & 1u) *reinterpret_cast<float*>(p + co[ii] + ii * 128) = a0;
            }
          }
          if ((sb1 | eb1) == 0u) {
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) a1 += y1[ii];
          } else {
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) {
              const float t = a1 + y1[ii];
              a1 = ((sb1 >> ii) & 1u) ? y1[ii] : t;
              if ((eb1 >> ii) & 1u) *reinterpret_cast<float*>(p + kAggRows * 128 + co[ii] + ii * 128) = a1;
            }
          }
        }
        // Sweep B: the warp that owns an end row stores its 512 B of segment sums (coalesced); the receiver id comes
        // from the owning lane's register
        named_bar_sync(1 + s, kTile);
        {
          const unsigned long long eh = q < 2 ? e0 : e1, fh = q < 2 ? f0 : f1;
          unsigned ends = (unsigned)(eh >> ((q & 1) * 32));
          // the first end of a 64-row block belongs to a segment that began in the previous block iff its flag bit 0 is clear
          const bool cont_first = !(fh & 1ull) && ((q & 1) == 0 || (unsigned)eh == 0u);
          const size_t agg_tile = (size_t)row0 / kAggRows + (q >> 1);
          bool first = true;
          while (ends != 0u) {
            const int i = __ffs((int)ends) - 1;
            ends &= ends - 1u;
            const int rid = __shfl_sync(0xffffffffu, cur_recv, i);
            const float4 v = *reinterpret_cast<const float4*>(S + (db[i & 7] - S_s) + (i & ~7) * 128);
            float* dst = (first && cont_first) ? b.carry + agg_tile * kH : b.aggpart + (size_t)rid * kH;
            *reinterpret_cast<float4*>(dst + 4 * lane) = v;
            first = false;
          }
        }
      }
      GM_ESTAMP(12);
      named_bar_sync(1 + s, kTile);            // every row has been read: the slot's staging tile is free
      GM_ESTAMP(13);
      if (has_next) begin_tile(k + KSTEP, false);
      GM_ESTAMP(14);
    }
  }
  tc_fence_before();
  cluster_sync_all();                          // the peer may still be arriving on / reading through this CTA
  if (warp == 0) tmem_dealloc_pair(tmem, 512);
}

// ================================================================================================
// mp_edge16_kernel: the same CTA-pair pipeline (memory plan, MMAs, barriers, fp32 order of the segmented sum) with
// EIGHT warps per 128-edge tile -- 544 threads per CTA: warps 0-7 slot 0, warps 8-15 slot 1, warp 16 the MMA issuer.
// Thread = edge row x two of the four 32-column boxes (warp w: rows 32 (w & 3) .., boxes h and h + 2 with
// h = (w >> 2) & 1, so every warp contributes to both K halves of the next operand and a K half is complete after
// ONE box of epilogue work per thread).  Sixteen compute warps at <= 120 registers (a thread holds 64 columns, not
// 128) instead of eight at 232: four warps per scheduler hide the latencies (tcgen05.ld, shared memory, L1 / L2)
// that bound the 8-warp kernel (ncu: issue slots 35 %, tensor pipe 33 %, nothing saturated), and the chain of a tile
// is half as long.  LayerNorm statistics of a row are combined from the two half rows (sum and centred sum of
// squares, exchanged through the row's own staging bytes); staging, residual reductions and the segment stores stay
// warp-local (a warp moves its own rows x its own boxes), the segmented sweep is thread = column x 64-row block.
// ================================================================================================
namespace {
constexpr int kNTE16 = 640;                  // 16 compute warps + the issuer's warpgroup (warp 16 issues, 17-19 idle)
constexpr int kIssuerWarp16 = 16;

// lane l of warp (q, h) moves float4 j16 of row r0 + 2 i + (l >> 4) in instruction i (16 instructions per 32 rows);
// float4 index in the row: box h + 2 ((l >> 3) & 1), chunk l & 7
__device__ __forceinline__ int stage16_j(int lane, int h) { return (h + 2 * ((lane >> 3) & 1)) * 8 + (lane & 7); }
// staging-tile offset of this lane's float4 of row r0 + 2 i + sub: (row & 7) = (2 i + sub) & 7 because r0 % 8 == 0
__device__ __forceinline__ uint32_t stage16_off(uint32_t db_base, uint32_t db_x, int i) {
  return db_base + (uint32_t)i * 256u + (db_x ^ (((uint32_t)i & 3u) << 5));
}
__device__ __forceinline__ void stage16_contig(uint32_t S_s, uint32_t db_base, uint32_t db_x, const float* src0,
                                               int n_valid, int sub) {
  if (n_valid >= 32) {
#pragma unroll
    for (int i = 0; i < 16; ++i) cp_async16(S_s + stage16_off(db_base, db_x, i), src0 + 2 * i * kH, 16u);
  } else {
#pragma unroll 4
    for (int i = 0; i < 16; ++i) {
      const bool ok = 2 * i + sub < n_valid;
      cp_async16(S_s + stage16_off(db_base, db_x, i), ok ? src0 + 2 * i * kH : src0, ok ? 16u : 0u);
    }
  }
  cp_async_commit();
}
__device__ __forceinline__ void stage16_gather(uint32_t S_s, uint32_t db_base, uint32_t db_x, const float* table_j,
                                               int my_idx, int sub) {
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int sid = __shfl_sync(0xffffffffu, my_idx, 2 * i + sub);
    cp_async16(S_s + stage16_off(db_base, db_x, i), table_j + (size_t)(sid < 0 ? 0 : sid) * kH, sid < 0 ? 0u : 16u);
  }
  cp_async_commit();
}
// ---- packed fp32 arithmetic (FADD2 / FMUL2 / FFMA2: two IEEE-rn operations per issued instruction; each half is
// bit-identical to the scalar op) for the elementwise parts of the epilogues ----
__device__ __forceinline__ float2 f2_add(float2 a, float2 b) {
  float2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<unsigned long long&>(a)), "l"(reinterpret_cast<unsigned long long&>(b)));
  return d;
}
__device__ __forceinline__ float2 f2_sub(float2 a, float2 b) {
  float2 d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<unsigned long long&>(a)), "l"(reinterpret_cast<unsigned long long&>(b)));
  return d;
}
__device__ __forceinline__ float2 f2_mul(float2 a, float2 b) {
  float2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<unsigned long long&>(a)), "l"(reinterpret_cast<unsigned long long&>(b)));
  return d;
}
__device__ __forceinline__ float2 f2_fma(float2 a, float2 b, float2 c) {
  float2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<unsigned long long&>(a)), "l"(reinterpret_cast<unsigned long long&>(b)),
        "l"(reinterpret_cast<unsigned long long&>(c)));
  return d;
}
__device__ __forceinline__ float2 f2_relu(float2 a) { return make_float2(fmaxf(a.x, 0.f), fmaxf(a.y, 0.f)); }
// bf16 hi / lo split of a pair (same values as split_bf16x2, the residual as one packed subtraction)
__device__ __forceinline__ void split_pair(float2 x, uint32_t& hi, uint32_t& lo) {
  __nv_bfloat162 h = __floats2bfloat162_rn(x.x, x.y);
  hi = *reinterpret_cast<uint32_t*>(&h);
  const float2 r = f2_sub(x, make_float2(__uint_as_float(hi << 16), __uint_as_float(hi & 0xFFFF0000u)));
  __nv_bfloat162 l = __floats2bfloat162_rn(r.x, r.y);
  lo = *reinterpret_cast<uint32_t*>(&l);
}
__device__ __forceinline__ void put_operand32_p(const float2* v, uint32_t tAH, uint32_t tAL, int c32) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) split_pair(v[j], hi[j], lo[j]);
  tmem_st16(tAH + c32 * 16, hi);
  tmem_st16(tAL + c32 * 16, lo);
}
__device__ __forceinline__ void put_operand32_inplace_p(const float2* v, uint32_t tR, int c32) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) split_pair(v[j], hi[j], lo[j]);
  tmem_st16(tR + c32 * 32, hi);
  tmem_st16(tR + c32 * 32 + 16, lo);
}
}  // namespace

__global__ void __launch_bounds__(kNTE16, 1) mp_edge16_kernel(TcMlp m, TcGraph g, TcBuffers b) {
  extern __shared__ __align__(1024) char smem_raw[];
  char* base = smem_raw;
  const uint32_t W_s = smem_u32(base);
  char* S_base = base + kWBytes;
  EdgeShared* sh = reinterpret_cast<EdgeShared*>(S_base + 2 * kStageBytes);
  // (the shuffle tells the compiler that the warp index is warp-uniform: role branches become uniform branches and the
  // uniform datapath stays usable inside them -- without it every global access re-creates its descriptor with R2UR)
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const uint32_t rank = cluster_ctarank();     // 0: leader (issues the pair's MMAs), 1: peer
  if (tid == 0 && (W_s & 1023u) != 0u) {
    printf("gnn_manip_b200: mp_edge16 dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  if (tid == 0) {
    mbar_init(smem_u32(&sh->w_full), 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&sh->a_half[s][0]), 16);   // 8 warps of the slot in each CTA of the pair
      mbar_init(smem_u32(&sh->a_half[s][1]), 16);
      mbar_init(smem_u32(&sh->d_ready[s]), 1);
    }
    fence_mbar_init();
    mbar_expect_tx(smem_u32(&sh->w_full), kWBytes);
    for (int l = 0; l < 3; ++l)
      for (int p = 0; p < 4; ++p)
        bulk_g2s(W_s + l * kWLayer + p * kWBlock, (const char*)m.wc[l] + (size_t)p * kPieceBytes + rank * kWBlock,
                 kWBlock, smem_u32(&sh->w_full));
  }
  if (warp == 0) tmem_alloc_pair(smem_u32(&sh->tmem_base), 512);
  for (int i = tid; i < kH; i += kNTE16) {
    sh->prm[kOffB2 + i] = m.b2[i];
    sh->prm[kOffB3 + i] = m.b3[i];
    sh->prm[kOffG + i] = m.ln_g[i];
    sh->prm[kOffBeta + i] = m.ln_b[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_base;
                                 // everything above touched constant parameters only

  const int E = min(*g.n_edges_dev, g.max_edges);
  const int n_tiles = (E + kTile - 1) / kTile;
  const int grid = (int)gridDim.x;
  const int b_lead = (int)blockIdx.x - (int)rank;
  const int n_my = b_lead < n_tiles ? (n_tiles - b_lead + grid - 1) / grid : 0;
  if (warp >= kIssuerWarp16) {
    // ---------------- fifth warpgroup: hands its registers to the compute warps (a scheduler holds four compute warps
    //                  and one warp of this group: 4 x 120 + 24 registers per lane); lane 0 of warp 16 in the leader
    //                  CTA issues every MMA: items of a slot in order (tile, layer, K half) ----------------
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    mbar_wait(smem_u32(&sh->w_full), 0);
    cluster_sync_all();                        // pairs with the compute warps' prologue barrier
    if (warp == kIssuerWarp16 && lane == 0 && rank == 0) {
      int it[2] = {0, 0};
      const int end[2] = {6 * ((n_my + 1) / 2), 6 * (n_my / 2)};
      uint32_t idle = 0;
      while (it[0] < end[0] || it[1] < end[1]) {
        bool progress = false;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          if (it[s] < end[s]) {
            const int half = it[s] & 1, layer = (it[s] >> 1) % 3;
            if (mbar_try_wait(smem_u32(&sh->a_half[s][half]), (uint32_t)(it[s] >> 1) & 1u)) {
              tc_fence_after();
              const uint32_t tXs = tmem + 256u * s, tYs = tXs + 128u;
              issue_half_pair(W_s + layer * kWLayer, layer == 1 ? tXs : tYs, layer == 1 ? tYs : tXs, half);
              if (half) umma_commit_pair(smem_u32(&sh->d_ready[s]));
              ++it[s];
              progress = true;
            }
          }
        }
        if (progress) idle = 0;
        else if (++idle > (1u << 28)) {
          printf("gnn_manip_b200: mp_edge16 issuer timed out (block %d)\n", blockIdx.x);
          __trap();
        }
      }
    }
    __syncwarp();
  } else {
    // ---------------- compute warps: slot s, thread = edge row r x boxes {h, h + 2} ----------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
    const int s = warp >> 3, q = warp & 3, h = (warp >> 2) & 1, r0 = q * 32, r = r0 + lane, sub = lane >> 4;
    const int c0 = h, c1 = h + 2;              // this thread's 32-column boxes: K half 0 and K half 1 of an operand
    const int KSTEP = 2, K0 = s;               // slot s takes this CTA's tiles s, s + 2, ...
    const int bar_slot = 1 + s, bar_rowpair = 3 + 4 * s + q;   // named barriers: the slot (256), the two warps of a row (64)
    char* S = S_base + s * kStageBytes;
    const uint32_t S_s = smem_u32(S);
    // row-owner offset of chunk j of a box: r * 128 + ((j ^ (r & 7)) << 4), rematerialised (one LOP3) at every use
    const uint32_t ro_base = (uint32_t)r * 128u, ro_x = ((uint32_t)r & 7u) << 4;
#define RO16(j) (ro_base | (ro_x ^ ((uint32_t)(j) << 4)))
    // regions of this slot: layer 1 reads X and accumulates into Y, layer 2 reads Y -> X, layer 3 reads X -> Y
    const uint32_t tX = tmem + 256u * s + ((uint32_t)r0 << 16), tY = tX + 128u;   // this warp's lanes
    const uint32_t bar_d = smem_u32(&sh->d_ready[s]);
    const uint32_t bar_a0 = mapa_shared(smem_u32(&sh->a_half[s][0]), 0);           // leader CTA's
    const uint32_t bar_a1 = mapa_shared(smem_u32(&sh->a_half[s][1]), 0);
    const float* prm = sh->prm;
    auto publish_half = [&](int half) {        // this warp's box of K half `half` of the next operand is in TMEM
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(half ? bar_a1 : bar_a0);
    };
    auto publish_both = [&]() {
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive_cluster(bar_a0);
        mbar_arrive_cluster(bar_a1);
      }
    };
    // ids of this thread's row in the slot's next tile; nx_send carries "this row starts a receiver segment" in bit 30
    // (an invalid row is -1 / -1: flag set, like the 4-warp kernel's -1 != -2)
    int nx_recv = -1, nx_send = -1;
    auto fetch_ids = [&](int k) {
      const int ge = (blockIdx.x + k * grid) * kTile + r;
      nx_recv = nx_send = -1;
      if (ge < E) {
        nx_recv = __ldg(g.erecv + ge);
        const int prev = ge > 0 ? __ldg(g.erecv + ge - 1) : -2;
        nx_send = __ldg(g.esend + ge) | (nx_recv != prev ? 1 << 30 : 0);
      }
    };
    // stager destination of this lane's float4 of row r0 + 2 i + sub: rematerialised, see stage16_off
    const int j16 = stage16_j(lane, h);
    const uint32_t db_base = (uint32_t)(j16 >> 3) * kBoxBytes + (uint32_t)(r0 + sub) * 128u;
    const uint32_t db_x = (((uint32_t)j16 & 7u) ^ (uint32_t)sub) << 4;
#define DB16(i) (db_base + (uint32_t)(i) * 256u + (db_x ^ (((uint32_t)(i) & 3u) << 5)))
    auto stage_e_rows = [&](int k) {           // this warp's 32 rows x 2 boxes of tile k
      const int ge0 = (blockIdx.x + k * grid) * kTile + r0;
      stage16_contig(S_s, db_base, db_x, b.e + (size_t)(ge0 + sub < E ? ge0 + sub : 0) * kH + 4 * j16, E - ge0, sub);
    };
    int cur_recv = -1;
    auto box_to_operand = [&](int c, uint32_t tR) {
      float2 v[16];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 t = *reinterpret_cast<const float4*>(S + c * kBoxBytes + RO16(j));
        v[2 * j] = make_float2(t.x, t.y); v[2 * j + 1] = make_float2(t.z, t.w);
      }
      put_operand32_inplace_p(v, tR, c);
    };
    // e rows of a tile (landed in this warp's staging bytes) -> bf16 hi / lo operand of layer 1 in TMEM
    auto operand_from_stage = [&]() {
      cp_async_wait_all();
      __syncwarp();
      box_to_operand(c0, tX);
      box_to_operand(c1, tX);
      tmem_st_wait();
      __syncwarp();                            // every lane is done reading this warp's staging bytes
    };
    auto begin_tile = [&]() {
      cur_recv = nx_recv;
      {
        const unsigned mk = __ballot_sync(0xffffffffu, (nx_send >> 30) & 1);
        if (lane == 0 && h == 0) sh->segmask[s][q] = mk;
      }
      const float* pr_row = b.pr + (size_t)max(cur_recv, 0) * kH;
      prefetch_l1(pr_row + 32 * c0);
      prefetch_l1(pr_row + 32 * c1);
      stage16_gather(S_s, db_base, db_x, b.ps + 4 * j16, nx_send < 0 ? -1 : (nx_send & 0x3fffffff), sub);
    };
    if (K0 < n_my) {
      fetch_ids(K0);
      stage_e_rows(K0);
    }
    // weights of this CTA landed; peer barriers initialised, peer weights landed, TMEM of the pair allocated
    mbar_wait(smem_u32(&sh->w_full), 0);
    cluster_sync_all();
    if (K0 < n_my) {
      operand_from_stage();
      publish_both();                            // layer 1 of the first tile needs e and weights only: its GEMM runs
                                                 // before the dependency wait as well
    }
    // Everything so far read only constants and e (written by the kernel two launches back); Pr / Ps come from the
    // node kernel right before.
    pdl_wait();
    pdl_launch_dependents();
    if (K0 < n_my) begin_tile();
    for (int k = K0; k < n_my; k += KSTEP) {
      const int row0 = (blockIdx.x + k * grid) * kTile;
      const int rows_valid = min(kTile, E - row0);
      const bool has_next = k + KSTEP < n_my;
#ifdef GM_STAMPS
      const bool stamp_on = blockIdx.x == 0 && tid == GM_STAMP_TID && k == 4;
#endif
      GM_ESTAMP(0);                              // stamps 0..14 in program order (tools_stamps.py, -DGM_STAMPS)
      if (has_next) fetch_ids(k + KSTEP);
      const uint32_t pt = (uint32_t)((k - K0) >> 1) & 1u;   // d_ready is used three times per tile: parities pt, pt ^ 1, pt
      const float* pr_row = b.pr + (size_t)max(cur_recv, 0) * kH;
      // ---- layer 1 epilogue: relu(D + Pr[recv] (b1 folded in) + Ps[send]) ----
      cp_async_wait_all();
      __syncwarp();
      GM_ESTAMP(1);
      mbar_wait(bar_d, pt);
      tc_fence_after();
      GM_ESTAMP(2);
      {
        float2 va[16];
        float4 p[8];
        auto load_pr = [&](int c) {
#pragma unroll
          for (int j = 0; j < 8; ++j) p[j] = __ldg(reinterpret_cast<const float4*>(pr_row + c * 32) + j);
        };
        auto e1_block = [&](int c, float2* v) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 s4 = *reinterpret_cast<const float4*>(S + c * kBoxBytes + RO16(j));
            v[2 * j] = f2_relu(f2_add(v[2 * j], f2_add(make_float2(p[j].x, p[j].y), make_float2(s4.x, s4.y))));
            v[2 * j + 1] = f2_relu(f2_add(v[2 * j + 1], f2_add(make_float2(p[j].z, p[j].w), make_float2(s4.z, s4.w))));
          }
          put_operand32_inplace_p(v, tY, c);
        };
        tmem_ld32(tY + 32 * c0, reinterpret_cast<float*>(va));
        load_pr(c0);
        tmem_ld_wait();
        e1_block(c0, va);
        tmem_ld32(tY + 32 * c1, reinterpret_cast<float*>(va));
        load_pr(c1);
        publish_half(0);
        tmem_ld_wait();
        e1_block(c1, va);
      }
      publish_half(1);
      GM_ESTAMP(3);
      // this warp's staging bytes are free until its LayerNorm rows go in: the e rows of the slot's next tile land
      // here while layers 2 and 3 run
      if (has_next) stage_e_rows(k + KSTEP);
      // ---- layer 2 epilogue: relu(D + b2) ----
      mbar_wait(bar_d, pt ^ 1u);
      tc_fence_after();
      GM_ESTAMP(4);
      {
        float2 va[16];
        auto e2_block = [&](int c, float2* v) {
          float4 bs[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) bs[j] = GM_PRM4(prm + kOffB2 + c * 32 + 4 * j);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            v[2 * j] = f2_relu(f2_add(v[2 * j], make_float2(bs[j].x, bs[j].y)));
            v[2 * j + 1] = f2_relu(f2_add(v[2 * j + 1], make_float2(bs[j].z, bs[j].w)));
          }
          put_operand32_inplace_p(v, tX, c);
        };
        tmem_ld32(tX + 32 * c0, reinterpret_cast<float*>(va));
        tmem_ld_wait();
        e2_block(c0, va);
        tmem_ld32(tX + 32 * c1, reinterpret_cast<float*>(va));
        publish_half(0);
        tmem_ld_wait();
        e2_block(c1, va);
      }
      publish_half(1);
      GM_ESTAMP(5);
      // ---- layer 3 epilogue: y = LayerNorm(D + b3) -> this thread's staging bytes ----
      mbar_wait(bar_d, pt);
      tc_fence_after();
      GM_ESTAMP(6);
      if (has_next) operand_from_stage();
      GM_ESTAMP(7);      // the A region is free (GEMM 3 is complete); D is read right below
      {
        // Two passes over the accumulator (tcgen05.ld twice; 32 live values instead of 64 keeps the kernel free of
        // spills at 120 registers).  Pass 1: per 32-column box an exact two-pass (sum, centred sum of squares), the
        // two boxes and then the two half rows combined like a two-pass variance (Chan et al.):
        // M2 = M2_a + M2_b + n_a (mean_a - mean)^2 + n_b (mean_b - mean)^2.
        float sum_h = 0.f, m2_h = 0.f, mean_c0 = 0.f;
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c = cc ? c1 : c0;
          float2 x[16];
          tmem_ld32(tY + 32 * c, reinterpret_cast<float*>(x));
          float4 t[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) t[j] = GM_PRM4(prm + kOffB3 + c * 32 + 4 * j);
          tmem_ld_wait();
          float2 sab = make_float2(0.f, 0.f), scd = sab;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            x[2 * j] = f2_add(x[2 * j], make_float2(t[j].x, t[j].y));
            x[2 * j + 1] = f2_add(x[2 * j + 1], make_float2(t[j].z, t[j].w));
            sab = f2_add(sab, x[2 * j]);
            scd = f2_add(scd, x[2 * j + 1]);
          }
          const float sum_c = (sab.x + sab.y) + (scd.x + scd.y), mean_c = sum_c * (1.0f / 32.0f);
          const float2 mc2 = make_float2(mean_c, mean_c);
          float2 qab = make_float2(0.f, 0.f), qcd = qab;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float2 u01 = f2_sub(x[2 * j], mc2), u23 = f2_sub(x[2 * j + 1], mc2);
            qab = f2_fma(u01, u01, qab);
            qcd = f2_fma(u23, u23, qcd);
          }
          const float m2_c = (qab.x + qab.y) + (qcd.x + qcd.y);
          if (cc == 0) {
            sum_h = sum_c; m2_h = m2_c; mean_c0 = mean_c;
          } else {
            const float dm = mean_c - mean_c0;
            sum_h += sum_c;
            m2_h = (m2_h + m2_c) + 16.0f * (dm * dm);   // n_a n_b / (n_a + n_b) = 16
          }
        }
        // exchange with the thread that owns the other two boxes of this row (warp w ^ 4): each side parks its pair
        // in its own staging bytes (free: the e rows were consumed above), reads the other's between two barriers
        GM_ESTAMP(8);
        *reinterpret_cast<float2*>(S + c0 * kBoxBytes + RO16(0)) = make_float2(sum_h, m2_h);
        named_bar_sync(bar_rowpair, 64);
        const float2 other = *reinterpret_cast<const float2*>(S + (c0 ^ 1) * kBoxBytes + RO16(0));
        named_bar_sync(bar_rowpair, 64);
        GM_ESTAMP(9);
        const float mean = (sum_h + other.x) * (1.0f / kH);
        const float dm = (sum_h - other.x) * (1.0f / 64.0f);           // mean_h - mean_other
        const float m2 = (m2_h + other.y) + 32.0f * (dm * dm);          // n_a n_b / (n_a + n_b) = 32
        const float rstd = 1.0f / sqrtf(m2 * (1.0f / kH) + 1e-5f);
        // Pass 2: normalise and park the row in this thread's staging bytes
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c = cc ? c1 : c0;
          float2 x[16];
          tmem_ld32(tY + 32 * c, reinterpret_cast<float*>(x));
          float4 t[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) t[j] = GM_PRM4(prm + kOffB3 + c * 32 + 4 * j);
          tmem_ld_wait();
          const float2 mean2 = make_float2(mean, mean), rstd2 = make_float2(rstd, rstd);
          // the accumulator has been read for the last time and the next tile's layer-1 operand is in TMEM: its
          // GEMM 1 runs under the residual / segmented-sum phases of this tile
          if (cc == 1 && has_next) publish_both();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 g4 = GM_PRM4(prm + kOffG + c * 32 + 4 * j);
            const float4 b4 = GM_PRM4(prm + kOffBeta + c * 32 + 4 * j);
            // ((x + b3) - mean) * rstd * g + beta, the last two as one fused multiply-add (as the scalar code compiles)
            const float2 y01 = f2_fma(f2_mul(f2_sub(f2_add(x[2 * j], make_float2(t[j].x, t[j].y)), mean2), rstd2),
                                      make_float2(g4.x, g4.y), make_float2(b4.x, b4.y));
            const float2 y23 = f2_fma(f2_mul(f2_sub(f2_add(x[2 * j + 1], make_float2(t[j].z, t[j].w)), mean2), rstd2),
                                      make_float2(g4.z, g4.w), make_float2(b4.z, b4.w));
            *reinterpret_cast<float4*>(S + c * kBoxBytes + RO16(j)) = make_float4(y01.x, y01.y, y23.x, y23.y);
          }
        }
      }
      __syncwarp();
      GM_ESTAMP(10);
      // ---- e' = e + y (epd_gnn.py:104): this warp's rows x boxes leave as vector reductions, the add happens in L2 ----
      {
        const int nr = min(32, rows_valid - r0);
        float* e_out = b.e + (size_t)(row0 + r0 + sub) * kH + 4 * j16;
        if (nr >= 32) {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            red_add_v4(e_out + 2 * i * kH, *reinterpret_cast<const float4*>(S + DB16(i)));
        } else {
#pragma unroll 4
          for (int i = 0; i < 16; ++i)
            if (2 * i + sub < nr)
              red_add_v4(e_out + 2 * i * kH, *reinterpret_cast<const float4*>(S + DB16(i)));
        }
      }
      GM_ESTAMP(11);
      named_bar_sync(bar_slot, 256);
      GM_ESTAMP(12);
      // ---- receiver-segmented sum of the pre-residual rows (scatter_sum, gn_utils.py:60-64): thread = column r of
      //      64-row block h.  Sweep A: running sum along the rows, restarted at segment starts and written back in
      //      place at segment ends (uniform bit masks, strictly sequential fp32 order).  Sweep B: one 512 B store per
      //      segment end -> aggpart[receiver] (or carry[] for a segment that began before this 64-row block). ----
      {
        const uint32_t cbase = (uint32_t)(r >> 5) * kBoxBytes + ((uint32_t)r & 3u) * 4u;
        const uint32_t cchunk = ((uint32_t)r >> 2) & 7u;
        uint32_t co[8];
#pragma unroll
        for (int ii = 0; ii < 8; ++ii) co[ii] = cbase + ((cchunk ^ (uint32_t)ii) << 4);
        const int n0 = max(0, min(kAggRows, rows_valid)), n1 = max(0, min(kAggRows, rows_valid - kAggRows));
        const unsigned long long f0 = ((unsigned long long)sh->segmask[s][1] << 32) | sh->segmask[s][0];
        const unsigned long long f1 = ((unsigned long long)sh->segmask[s][3] << 32) | sh->segmask[s][2];
        const unsigned long long v0 = n0 >= 64 ? ~0ull : ((1ull << n0) - 1ull), v1 = n1 >= 64 ? ~0ull : ((1ull << n1) - 1ull);
        const unsigned long long m0 = f0 & ~1ull & v0, m1 = f1 & ~1ull & v1;              // restart points
        const unsigned long long e0 = ((m0 >> 1) | (n0 > 0 ? 1ull << (n0 - 1) : 0ull)) & v0;   // segment ends
        const unsigned long long e1 = ((m1 >> 1) | (n1 > 0 ? 1ull << (n1 - 1) : 0ull)) & v1;
        {
          const unsigned long long mh = h ? m1 : m0, eh = h ? e1 : e0;
          float a0 = 0.f;
#pragma unroll 1
          for (int bb = 0; bb < kAggRows / 8; ++bb) {
            float y0[8];
            char* p = S + h * (kAggRows * 128) + bb * 1024;
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) y0[ii] = *reinterpret_cast<const float*>(p + co[ii] + ii * 128);
            const unsigned sb0 = (unsigned)(mh >> (8 * bb)) & 0xffu, eb0 = (unsigned)(eh >> (8 * bb)) & 0xffu;
            if ((sb0 | eb0) == 0u) {
#pragma unroll
              for (int ii = 0; ii < 8; ++ii) a0 += y0[ii];
            } else {
#pragma unroll
              for (int ii = 0; ii < 8; ++ii) {
                const float t = a0 + y0[ii];
                a0 = ((sb0 >> ii) & 1u) ? y0[ii] : t;
                if ((eb0 >> ii) & 1u) *reinterpret_cast<float*>(p + co[ii] + ii * 128) = a0;
              }
            }
          }
        }
        named_bar_sync(bar_slot, 256);
        {
          // warp (q, h) owns the segment ends of rows 32 q + 16 h .. + 15 (64-row block q >> 1)
          const unsigned long long eb = q < 2 ? e0 : e1, fb = q < 2 ? f0 : f1;
          const int off = (q & 1) * 32 + 16 * h;
          unsigned ends = (unsigned)(eb >> off) & 0xffffu;
          // the first end of a 64-row block belongs to a segment that began in the previous block iff flag bit 0 is clear
          const bool cont_first = !(fb & 1ull) && (eb & ((1ull << off) - 1ull)) == 0ull;
          const size_t agg_tile = (size_t)row0 / kAggRows + (q >> 1);
          bool first = true;
          while (ends != 0u) {
            const int i = 16 * h + __ffs((int)ends) - 1;
            ends &= ends - 1u;
            const int rid = __shfl_sync(0xffffffffu, cur_recv, i);
            const float4 v = *reinterpret_cast<const float4*>(S + stage_chunk_off(r0 + i, lane));
            float* dst = (first && cont_first) ? b.carry + agg_tile * kH : b.aggpart + (size_t)rid * kH;
            *reinterpret_cast<float4*>(dst + 4 * lane) = v;
            first = false;
          }
        }
      }
      GM_ESTAMP(13);
      named_bar_sync(bar_slot, 256);           // every row has been read: the slot's staging tile is free
      GM_ESTAMP(14);
      if (has_next) begin_tile();
      GM_ESTAMP(15);
    }
  }
  tc_fence_before();
  cluster_sync_all();                          // the peer may still be arriving on / reading through this CTA
  if (warp == 0) tmem_dealloc_pair(tmem, 512);
}
#undef RO16
#undef DB16

// ================================================================================================
// enc_edge (pair version): edge encoder of EncProcDecGNN.forward (epd_gnn.py:30-33,88) on the same CTA-pair
// pipeline.  Edge features (row A2, utils.py:43-61) are built on the fly from positions; Linear(4,128)+ReLU runs
// on CUDA cores straight into the TMEM operand, the two 128x128 layers on the tensor cores with both weight chunks
// resident, LayerNorm in registers, e rows leave through this warp's staging rows as coalesced 512 B stores.
// No cross-warp hand-off exists inside a CTA besides the MMA barriers.
// ================================================================================================
__global__ void __launch_bounds__(kNTE, 1)
enc_edge_pair_kernel(TcMlp m, const float* __restrict__ pos, int pos_stride, float conn_r,
                     const float* __restrict__ edge_attr, const int* __restrict__ eref, TcGraph g, TcBuffers b) {
  extern __shared__ __align__(1024) char smem_raw[];
  char* base = smem_raw;
  const uint32_t W_s = smem_u32(base);
  // first layer (4 -> 128), stored for PAIRS of outputs (a, b) = (2p, 2p + 1): w1p[8 p ..] = {wa.x, wb.x, wa.y, wb.y, wa.z,
  // wb.z, wa.w, wb.w}, so that two float4 loads give the operands of four packed FMAs (f2_fma) for both outputs
  float* w1p = reinterpret_cast<float*>(base + 2 * kWLayer);             // [64][8]
  float* b1s = reinterpret_cast<float*>(base + 2 * kWLayer + kH * 16);   // [128]
  char* S_base = base + kWBytes;
  EdgeShared* sh = reinterpret_cast<EdgeShared*>(S_base + 2 * kStageBytes);
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;   // (warp-uniform)
  const uint32_t rank = cluster_ctarank();
  if (tid == 0) {
    mbar_init(smem_u32(&sh->w_full), 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&sh->a_ready[s]), 8);
      mbar_init(smem_u32(&sh->d_ready[s]), 1);
    }
    fence_mbar_init();
    mbar_expect_tx(smem_u32(&sh->w_full), 2 * kWLayer);
    for (int l = 0; l < 2; ++l)
      for (int p = 0; p < 4; ++p)
        bulk_g2s(W_s + l * kWLayer + p * kWBlock, (const char*)m.wc[2 + l] + (size_t)p * kPieceBytes + rank * kWBlock,
                 kWBlock, smem_u32(&sh->w_full));
  }
  if (warp == 0) tmem_alloc_pair(smem_u32(&sh->tmem_base), 512);
  for (int i = tid; i < kH; i += kNTE) {
    sh->prm[kOffB2 + i] = m.b2[i];
    sh->prm[kOffB3 + i] = m.b3[i];
    sh->prm[kOffG + i] = m.ln_g[i];
    sh->prm[kOffBeta + i] = m.ln_b[i];
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) w1p[(i >> 1) * 8 + 2 * k4 + (i & 1)] = m.w1[4 * i + k4];
    b1s[i] = m.b1[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_base;
  pdl_wait();
  pdl_launch_dependents();

  const int E = min(*g.n_edges_dev, g.max_edges);
  const int n_tiles = (E + kTile - 1) / kTile;
  const int grid = (int)gridDim.x;
  const int b_lead = (int)blockIdx.x - (int)rank;
  const int n_my = b_lead < n_tiles ? (n_tiles - b_lead + grid - 1) / grid : 0;
  const int s = warp >> 2, q = warp & 3, r = tid & (kTile - 1), r0 = q * 32;
  char* S = S_base + s * kStageBytes;
  const uint32_t S_s = smem_u32(S);
  uint32_t ro[8], db[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) ro[j] = stage_chunk_off(r, j);
  stage_dst_init(S_s, r0, lane, db);
  const uint32_t tD = tmem + ((uint32_t)r0 << 16) + 256u * s, tAH = tD + 128u, tAL = tD + 192u;
  const uint32_t bar_a_local = smem_u32(&sh->a_ready[s]), bar_d = smem_u32(&sh->d_ready[s]);
  const uint32_t bar_a = mapa_shared(bar_a_local, 0);
  const bool issuer = rank == 0 && q == 0 && lane == 0;
  const float* prm = sh->prm;
  uint32_t ph_d = 0, ph_a = 0;
  int layer = 0;
  auto issue_gemm = [&]() {
    if (issuer) {
      mbar_wait(bar_a_local, ph_a);
      tc_fence_after();
      issue_layer_pair(W_s + layer * kWLayer, tmem + 256u * s, tmem + 256u * s + 128u, tmem + 256u * s + 192u);
      umma_commit_pair(bar_d);
    }
    ph_a ^= 1u;
    layer ^= 1;
    __syncwarp();
  };
  auto load_x = [&](int k) -> float4 {
    const int ge = (blockIdx.x + k * grid) * kTile + r;
    if (ge >= E) return make_float4(0.f, 0.f, 0.f, 0.f);
    if (edge_attr != nullptr)      // explicit [E, 4] features in the caller's edge order (eref: receiver-major -> caller's id)
      return __ldg(reinterpret_cast<const float4*>(edge_attr) + (eref ? __ldg(eref + ge) : ge));
    return edge_feature(pos, pos_stride, __ldg(g.esend + ge), __ldg(g.erecv + ge), conn_r);
  };
  // layer 1 on CUDA cores: relu(W1 x + b1) -> bf16 hi / lo operand in TMEM (same FMA order as the SIMT path)
  auto layer1 = [&](const float4& x) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float2 v[16];
      const float2 xx = make_float2(x.x, x.x), xy = make_float2(x.y, x.y), xz = make_float2(x.z, x.z), xw = make_float2(x.w, x.w);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float4 w01 = *reinterpret_cast<const float4*>(w1p + (c * 16 + j) * 8);
        const float4 w23 = *reinterpret_cast<const float4*>(w1p + (c * 16 + j) * 8 + 4);
        const float2 bb = *reinterpret_cast<const float2*>(b1s + c * 32 + 2 * j);
        // per output: fma(x.w, w.w, fma(x.z, w.z, fma(x.y, w.y, fma(x.x, w.x, b)))) -- the SIMT path's order
        v[j] = f2_relu(f2_fma(xw, make_float2(w23.z, w23.w), f2_fma(xz, make_float2(w23.x, w23.y),
                       f2_fma(xy, make_float2(w01.z, w01.w), f2_fma(xx, make_float2(w01.x, w01.y), bb)))));
      }
      put_operand32_p(v, tAH, tAL, c);
    }
  };
  float4 xn = make_float4(0.f, 0.f, 0.f, 0.f);
  if (s < n_my) xn = load_x(s);
  mbar_wait(smem_u32(&sh->w_full), 0);
  cluster_sync_all();
  if (s < n_my) {                                  // first tile of the slot: operand and GEMM 2 up front
    layer1(xn);
    publish_operand(bar_a, lane);
    issue_gemm();
    if (s + 2 < n_my) xn = load_x(s + 2);
  }
  for (int k = s; k < n_my; k += 2) {
    const int row0 = (blockIdx.x + k * grid) * kTile;
    const int rows_valid = min(kTile, E - row0);
    const bool has_next = k + 2 < n_my;
    const float4 x_next = xn;                      // features of this slot's next tile (loaded during the previous one)
    if (k + 4 < n_my) xn = load_x(k + 4);          // in flight during this tile
    // ---- layer 2 epilogue: relu(D + b2) ----
    mbar_wait(bar_d, ph_d);
    ph_d ^= 1u;
    tc_fence_after();
    {
      float2 va[16], vb[16];
      auto e2_block = [&](int c, float2* v) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 bs = GM_PRM4(prm + kOffB2 + c * 32 + 4 * j);
          v[2 * j] = f2_relu(f2_add(v[2 * j], make_float2(bs.x, bs.y)));
          v[2 * j + 1] = f2_relu(f2_add(v[2 * j + 1], make_float2(bs.z, bs.w)));
        }
        put_operand32_p(v, tAH, tAL, c);
      };
      tmem_ld32(tD, reinterpret_cast<float*>(va));
      tmem_ld_wait();
      tmem_ld32(tD + 32, reinterpret_cast<float*>(vb));
      e2_block(0, va);
      tmem_ld_wait();
      tmem_ld32(tD + 64, reinterpret_cast<float*>(va));
      e2_block(1, vb);
      tmem_ld_wait();
      tmem_ld32(tD + 96, reinterpret_cast<float*>(vb));
      e2_block(2, va);
      tmem_ld_wait();
      e2_block(3, vb);
    }
    publish_operand(bar_a, lane);
    issue_gemm();
    // ---- layer 3 epilogue: e0 = LayerNorm(D + b3) -> this thread's staging row ----
    mbar_wait(bar_d, ph_d);
    ph_d ^= 1u;
    tc_fence_after();
    // the operand region is free (GEMM 3 is complete): the next tile's layer 1 goes in now, and its GEMM 2 runs
    // under this tile's LayerNorm and row stores once every warp has the accumulator in registers
    if (has_next) layer1(x_next);
    {
      // exact two-pass LayerNorm of the whole row in registers, packed fp32 arithmetic (same operations per element)
      float2 y[kH / 2];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tD + c * 32, reinterpret_cast<float*>(y) + c * 32);
      tmem_ld_wait();
      if (has_next) publish_operand(bar_a, lane);
      float2 sab = make_float2(0.f, 0.f), scd = sab;
#pragma unroll
      for (int j = 0; j < kH / 4; ++j) {
        const float4 t = GM_PRM4(prm + kOffB3 + 4 * j);
        y[2 * j] = f2_add(y[2 * j], make_float2(t.x, t.y));
        y[2 * j + 1] = f2_add(y[2 * j + 1], make_float2(t.z, t.w));
        sab = f2_add(sab, y[2 * j]);
        scd = f2_add(scd, y[2 * j + 1]);
      }
      const float mean = ((sab.x + sab.y) + (scd.x + scd.y)) * (1.0f / kH);
      const float2 mean2 = make_float2(mean, mean);
      float2 qab = make_float2(0.f, 0.f), qcd = qab;
#pragma unroll
      for (int j = 0; j < kH / 4; ++j) {
        y[2 * j] = f2_sub(y[2 * j], mean2);
        y[2 * j + 1] = f2_sub(y[2 * j + 1], mean2);
        qab = f2_fma(y[2 * j], y[2 * j], qab);
        qcd = f2_fma(y[2 * j + 1], y[2 * j + 1], qcd);
      }
      const float rstd = 1.0f / sqrtf(((qab.x + qab.y) + (qcd.x + qcd.y)) * (1.0f / kH) + 1e-5f);
      const float2 rstd2 = make_float2(rstd, rstd);
#pragma unroll
      for (int j = 0; j < kH / 4; ++j) {
        const float4 g4 = GM_PRM4(prm + kOffG + 4 * j);
        const float4 b4 = GM_PRM4(prm + kOffBeta + 4 * j);
        const float2 o01 = f2_fma(f2_mul(y[2 * j], rstd2), make_float2(g4.x, g4.y), make_float2(b4.x, b4.y));
        const float2 o23 = f2_fma(f2_mul(y[2 * j + 1], rstd2), make_float2(g4.z, g4.w), make_float2(b4.z, b4.w));
        *reinterpret_cast<float4*>(S + (j >> 3) * kBoxBytes + ro[j & 7]) = make_float4(o01.x, o01.y, o23.x, o23.y);
      }
    }
    __syncwarp();
    if (has_next) issue_gemm();                    // every warp of the pair has published by now
    // ---- this warp's 32 rows leave as coalesced 512 B stores ----
    {
      const int nr = min(32, rows_valid - r0);
      float* e_out = b.e + (size_t)(row0 + r0) * kH + 4 * lane;
      if (nr >= 32) {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          *reinterpret_cast<float4*>(e_out + i * kH) = *reinterpret_cast<const float4*>(S + (db[i & 7] - S_s) + (i & ~7) * 128);
      } else {
#pragma unroll 8
        for (int i = 0; i < 32; ++i)
          if (i < nr)
            *reinterpret_cast<float4*>(e_out + i * kH) =
                *reinterpret_cast<const float4*>(S + (db[i & 7] - S_s) + (i & ~7) * 128);
      }
    }
    __syncwarp();                                  // the rows are rewritten by this warp's next tile
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc_pair(tmem, 512);
}

// ---- self-test GEMM through the TMEM-operand path: out += x @ W^T (out must be zeroed by the caller) ----
__global__ void __launch_bounds__(128, 1)
selftest_gemm_ts_kernel(const float* __restrict__ x, const void* chunk, float* __restrict__ out, int M) {
  extern __shared__ char smem_raw[];
  char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t W_s = smem_u32(base);
  char* S = base + kChunkBytes;
  EdgeShared* sh = reinterpret_cast<EdgeShared*>(S + kStageBytes);
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;   // (warp-uniform)
  const uint32_t bar_w = smem_u32(&sh->w_full), bar_d = smem_u32(&sh->d_ready[0]);
  if (tid == 0) {
    mbar_init(bar_w, 1);
    mbar_init(bar_d, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&sh->tmem_base), 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_base;
  if (tid == 0) {
    mbar_expect_tx(bar_w, kChunkBytes);
    for (int i = 0; i < 4; ++i)
      bulk_g2s(W_s + i * kPieceBytes, (const char*)chunk + (size_t)i * kPieceBytes, kPieceBytes, bar_w);
  }
  const uint32_t S_s = smem_u32(S);
  const int r0 = warp * 32;
  uint32_t ro[8], db[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) ro[j] = stage_chunk_off(tid, j);
  stage_dst_init(S_s, r0, lane, db);
  const uint32_t tD = tmem + ((uint32_t)r0 << 16), tAH = tD + 128u, tAL = tD + 192u;
  uint32_t ph = 0;
  const int n_tiles = M / kTile;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const size_t row0 = (size_t)tile * kTile;
    stage_rows32_contig(db, x + (row0 + r0) * kH + 4 * lane, 32);
    cp_async_wait_all();
    __syncwarp();
    row_to_operand(S, ro, tAH, tAL);
    tmem_st_wait();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      mbar_wait(bar_w, 0);
      tc_fence_after();
      for (int p = 0; p < 4; ++p) issue_piece(p, W_s + p * kPieceBytes, tmem, tmem + 128u, tmem + 192u);
      umma_commit(bar_d);
    }
    __syncwarp();
    mbar_wait(bar_d, ph);
    ph ^= 1u;
    tc_fence_after();
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float v[32];
      tmem_ld32(tD + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(S + c * kBoxBytes + ro[j]) =
            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
    __syncwarp();
    for (int i = 0; i < 32; ++i) {
      const float4 y = *reinterpret_cast<const float4*>(S + stage_chunk_off(r0 + i, lane));
      red_add_v4(out + (row0 + r0 + i) * kH + 4 * lane, y);
    }
    tc_fence_before();
    __syncthreads();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

namespace {
constexpr int kSelftestTsSmem = kChunkBytes + kStageBytes + (int)sizeof(EdgeShared) + 1024;
}

int tc_debug_stamps_edge(unsigned long long* out_host, int n) {
#ifdef GM_STAMPS
  GM_CUDA(cudaMemcpyFromSymbol(out_host, g_edge_stamps, sizeof(unsigned long long) * (n < 32 ? n : 32)));
#else
  (void)out_host; (void)n;
#endif
  return GM_OK;
}

int tc_configure_edge() {
  static PerDeviceOnce once;
  const int dev = once.begin();
  if (dev == -1) return GM_OK;
  GM_REQUIRE(dev >= 0, "tc_configure_edge: cannot query the current device");
  GM_CUDA(cudaFuncSetAttribute(mp_edge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kEdgeSmemBytes));
  GM_CUDA(cudaFuncSetAttribute(mp_edge16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kEdgeSmemBytes));
  GM_CUDA(cudaFuncSetAttribute(enc_edge_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kEdgeSmemBytes));
  GM_CUDA(cudaFuncSetAttribute(selftest_gemm_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSelftestTsSmem));
  once.finish(dev);
  return GM_OK;
}

int tc_mp_edge(const TcMlp& m, const TcGraph& g, const TcBuffers& b, cudaStream_t stream, bool programmatic) {
  if (g.max_edges <= 0) return GM_OK;
  int rc = tc_configure_edge();
  if (rc) return rc;
  // CTA pairs (clusters of 2): even grid, at most one CTA per SM
  const int grid = min((ceil_div(g.max_edges, kTile) + 1) & ~1, kNumSMs & ~1);
  // programmatic = false: fully ordered after everything before it in the stream (its early part reads e, which
  // must then come from the kernel right before it or from a joined stream)
  // default: eight warps per tile (mp_edge16_kernel); GM_EDGE_KERNEL=8 keeps the 4-warps-per-tile kernel for A/B runs
  static const bool edge8 = [] { const char* e = getenv("GM_EDGE_KERNEL"); return e && e[0] == '8'; }();
  if (edge8) {
    GM_CUDA(launch_ex(mp_edge_kernel, grid, kNTE3, (size_t)kEdgeSmemBytes, stream, 2, programmatic, m, g, b));
  } else {
    GM_CUDA(launch_ex(mp_edge16_kernel, grid, kNTE16, (size_t)kEdgeSmemBytes, stream, 2, programmatic, m, g, b));
  }
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("mp_edge_kernel");
  return GM_OK;
}

// edge_dim == 4 (on-the-fly displacement features from positions, or an explicit [E, 4] edge_attr): the CTA-pair pipeline
int tc_enc_edge_pair(const TcMlp& m, const float* pos, int pos_stride, float conn_r, const float* edge_attr,
                     const int* eref, const TcGraph& g, const TcBuffers& b, cudaStream_t stream) {
  if (g.max_edges <= 0) return GM_OK;
  int rc = tc_configure_edge();
  if (rc) return rc;
  const int grid = min((ceil_div(g.max_edges, kTile) + 1) & ~1, kNumSMs & ~1);
  GM_CUDA(launch_pdl(enc_edge_pair_kernel, grid, kNTE, (size_t)kEdgeSmemBytes, stream, 2, m, pos, pos_stride, conn_r,
                     edge_attr, eref, g, b));
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("enc_edge_pair_kernel");
  return GM_OK;
}

int tc_selftest_gemm_ts(const float* x, const void* chunk, float* out, int M, cudaStream_t stream) {
  GM_REQUIRE(M > 0 && M % kTile == 0, "tc_selftest_gemm_ts: M=%d must be a positive multiple of %d", M, kTile);
  int rc = tc_configure_edge();
  if (rc) return rc;
  selftest_gemm_ts_kernel<<<min(M / kTile, kNumSMs), 128, kSelftestTsSmem, stream>>>(x, chunk, out, M);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("selftest_gemm_ts_kernel");
  return GM_OK;
}

GM_TRACE_DUMP(trace_dump_edge)

}  // namespace gm
