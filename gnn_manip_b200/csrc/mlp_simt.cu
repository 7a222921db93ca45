// Generic fp32 building blocks for the graph-network block API (any widths / depths):
//   gather + concat + MLP (+LayerNorm, +residual)      == EdgeModel / NodeModel phi call
//                                                         (deps/torch-graphnet/torch_graphnet/graph_networks.py:43-50, 75-85)
//   deterministic receiver-sorted segmented sum         == scatter_sum (utils/gn_utils.py:60-64)
// These are plain CUDA-core kernels: they serve the drop-in block classes for shapes the tcgen05 path
// does not cover, and are the on-device fp32 cross-check of the tensor-core kernels in mp_tc.cu.
#include "common.cuh"
#include "mlp_simt.cuh"

namespace gm {

namespace {

constexpr int TM = 32;        // rows per CTA
constexpr int NT = 128;       // threads per CTA

__global__ void __launch_bounds__(NT) mlp_simt_kernel(MlpSimtArgs a) {
  extern __shared__ __align__(16) float smem[];
  const int M = a.M_dev ? min(*a.M_dev, a.M) : a.M;
  const int row0 = blockIdx.x * TM;
  if (row0 >= M) return;
  const int rows = min(TM, M - row0);
  const int ld = a.max_width;           // leading dimension of both activation buffers (multiple of 4)
  float* act0 = smem;
  float* act1 = smem + (size_t)TM * ld;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- assemble the concatenated input rows (warp per row, coalesced over columns) ----
  for (int r = warp; r < TM; r += NT / 32) {
    float* dst = act0 + (size_t)r * ld;
    int col = 0;
    for (int s = 0; s < a.n_src; ++s) {
      const int w = a.width[s];
      if (r < rows) {
        const int m = row0 + r;
        const long long srow = a.idx[s] ? (long long)a.idx[s][m] : (long long)m;
        const float* src = a.src[s] + srow * a.src_stride[s];
        for (int c = lane; c < w; c += 32) dst[col + c] = src[c];
      } else {
        for (int c = lane; c < w; c += 32) dst[col + c] = 0.f;
      }
      col += w;
    }
    for (int c = col + lane; c < ld; c += 32) dst[c] = 0.f;
  }
  __syncthreads();

  float* in = act0;
  float* out = act1;
  for (int l = 0; l < a.n_layers; ++l) {
    const int K = a.layers[l].in, N = a.layers[l].out;
    const float* __restrict__ Wt = a.layers[l].Wt;
    const float* __restrict__ b = a.layers[l].b;
    const bool relu = l + 1 < a.n_layers;
    const int K4 = K & ~3;
    for (int c = tid; c < N; c += NT) {
      float acc[TM];
      const float bias = b ? b[c] : 0.f;
#pragma unroll
      for (int r = 0; r < TM; ++r) acc[r] = bias;
      for (int k = 0; k < K4; k += 4) {
        const float w0 = Wt[(size_t)k * N + c], w1 = Wt[(size_t)(k + 1) * N + c];
        const float w2 = Wt[(size_t)(k + 2) * N + c], w3 = Wt[(size_t)(k + 3) * N + c];
#pragma unroll
        for (int r = 0; r < TM; ++r) {
          const float4 x = *reinterpret_cast<const float4*>(in + (size_t)r * ld + k);
          acc[r] = fmaf(x.x, w0, acc[r]);
          acc[r] = fmaf(x.y, w1, acc[r]);
          acc[r] = fmaf(x.z, w2, acc[r]);
          acc[r] = fmaf(x.w, w3, acc[r]);
        }
      }
      for (int k = K4; k < K; ++k) {
        const float w0 = Wt[(size_t)k * N + c];
#pragma unroll
        for (int r = 0; r < TM; ++r) acc[r] = fmaf(in[(size_t)r * ld + k], w0, acc[r]);
      }
#pragma unroll
      for (int r = 0; r < TM; ++r) out[(size_t)r * ld + c] = relu ? fmaxf(acc[r], 0.f) : acc[r];
    }
    __syncthreads();
    float* t = in; in = out; out = t;
  }
  // `in` now holds the last layer's output [TM][Nout]
  const int Nout = a.layers[a.n_layers - 1].out;
  for (int r = warp; r < rows; r += NT / 32) {
    float* x = in + (size_t)r * ld;
    const long long m = row0 + r;
    if (a.ln_g) {
      float s = 0.f;
      for (int c = lane; c < Nout; c += 32) s += x[c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const float mean = s / (float)Nout;
      float v = 0.f;
      for (int c = lane; c < Nout; c += 32) { const float d = x[c] - mean; v = fmaf(d, d, v); }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      const float rstd = 1.0f / sqrtf(v / (float)Nout + a.ln_eps);
      for (int c = lane; c < Nout; c += 32) x[c] = (x[c] - mean) * rstd * a.ln_g[c] + a.ln_b[c];
    }
    const long long orow = a.out_perm ? (long long)a.out_perm[m] : m;
    for (int c = lane; c < Nout; c += 32) {
      const float y = x[c];
      if (a.out_pre) a.out_pre[orow * Nout + c] = y;
      if (a.out) a.out[orow * Nout + c] = a.residual ? y + a.residual[orow * Nout + c] : y;
    }
  }
}

// agg[n][:] = sum over the receiver segment of n, visiting edges in ascending reference edge id
// (bit-identical to the sequential index_add_ of the CPU reference).  Warp per node.
__global__ void segment_sum_kernel(const float* __restrict__ src, int F, const int* __restrict__ recv_ptr,
                                   const int* __restrict__ eref, int n_nodes, float* __restrict__ agg) {
  const int node = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (node >= n_nodes) return;
  const int b = recv_ptr[node], e = recv_ptr[node + 1];
  for (int c0 = 0; c0 < F; c0 += 128) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int t = b; t < e; ++t) {
      const float* row = src + (size_t)(eref ? eref[t] : t) * F + c0;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c = lane + 32 * i;
        if (c0 + c < F) acc[i] += row[c];
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = lane + 32 * i;
      if (c0 + c < F) agg[(size_t)node * F + c0 + c] = acc[i];
    }
  }
}

__global__ void gather_rows_kernel(const float* __restrict__ src, int F, const int* __restrict__ idx, long long M,
                                   float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M * F) return;
  const long long m = i / F;
  const int c = (int)(i - m * F);
  out[i] = src[(size_t)idx[m] * F + c];
}

}  // namespace

// opt in to > 48 KB dynamic shared memory once per device (not legal inside a stream capture)
int mlp_simt_configure() {
  static PerDeviceOnce once;
  const int dev = once.begin();
  if (dev == -1) return GM_OK;
  GM_REQUIRE(dev >= 0, "mlp_simt_configure: cannot query the current device");
  GM_CUDA(cudaFuncSetAttribute(mlp_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024)));
  once.finish(dev);
  return GM_OK;
}

int mlp_simt(const MlpSimtArgs& args, cudaStream_t stream) {
  GM_REQUIRE(args.n_layers >= 1 && args.n_layers <= kMaxSimtLayers, "mlp_simt: n_layers=%d", args.n_layers);
  GM_REQUIRE(args.n_src >= 1 && args.n_src <= 3, "mlp_simt: n_src=%d", args.n_src);
  if (args.M <= 0) return GM_OK;
  int in_w = 0;
  for (int s = 0; s < args.n_src; ++s) in_w += args.width[s];
  GM_REQUIRE(in_w == args.layers[0].in, "mlp_simt: concatenated input width %d != first layer in %d", in_w,
             args.layers[0].in);
  int mw = in_w;
  for (int l = 0; l < args.n_layers; ++l) {
    if (l > 0) GM_REQUIRE(args.layers[l].in == args.layers[l - 1].out, "mlp_simt: layer %d width mismatch", l);
    mw = mw > args.layers[l].out ? mw : args.layers[l].out;
  }
  MlpSimtArgs a = args;
  a.max_width = (int)align_up((size_t)mw, 4);
  const size_t smem = 2ull * TM * a.max_width * sizeof(float);
  GM_REQUIRE(smem <= 200 * 1024, "mlp_simt: layer width %d too large for shared memory", mw);
  int rc = mlp_simt_configure();
  if (rc) return rc;
  mlp_simt_kernel<<<ceil_div(a.M, TM), NT, smem, stream>>>(a);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("mlp_simt_kernel");
  return GM_OK;
}

int segment_sum(const float* src, int F, const int* recv_ptr, const int* eref, int n_nodes, float* agg,
                cudaStream_t stream) {
  if (n_nodes <= 0) return GM_OK;
  segment_sum_kernel<<<ceil_div(n_nodes, 4), 128, 0, stream>>>(src, F, recv_ptr, eref, n_nodes, agg);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("segment_sum_kernel");
  return GM_OK;
}

int gather_rows(const float* src, int F, const int* idx, long long M, float* out, cudaStream_t stream) {
  if (M <= 0) return GM_OK;
  gather_rows_kernel<<<ceil_div(M * F, 256), 256, 0, stream>>>(src, F, idx, M, out);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("gather_rows_kernel");
  return GM_OK;
}

}  // namespace gm
