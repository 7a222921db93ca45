// Row A1 -- capped radius graph as a cell-list spatial hash with a warp-cooperative neighbour scan.
//
// Replaces gnn_manip/utils/utils.py:64-93 (get_connectivity: sklearn KDTree.query_radius sorted by
// distance, first max_neighbours kept per query).  Truncation rule reproduced exactly:
//   for every query i (= sender): all j with d2(i,j) <= r*r (self included), ordered by ascending
//   distance (ties: ascending j -- undefined in the reference), first K kept.
// d2 is evaluated like scikit-learn's KDTree64: float32 inputs widened to float64,
// d2 = ((dx*dx) + dy*dy) + dz*dz with round-to-nearest and no FMA contraction
// (sklearn/metrics/_dist_metrics.pxd.tp:39-49, neighbors/_binary_tree.pxi.tp:1952-1967).
//
// Outputs (all int32, device):
//   nbr [n, K], cnt [n]     sender-major neighbour table == the reference edge list when flattened
//   edge_ptr [n+1]          exclusive scan of cnt: reference edge id of (i, s) is edge_ptr[i] + s
//   recv_ptr [n+1]          receiver CSR offsets
//   esend/erecv/eref [n*K]  edges sorted by (receiver, sender): sender, receiver, reference edge id
//   n_edges (device scalar) = edge_ptr[n]
// The receiver-major order is what the message-passing kernels consume: a receiver's incoming edges are
// contiguous, so the scatter_add of graph_networks.py:75-85 becomes a deterministic segmented sum that
// visits the edges of each receiver in ascending reference edge id (the CPU order of index_add_).
#include <stdlib.h>
#include "common.cuh"
#include "scan.cuh"
#include "radius_graph.cuh"

namespace gm {

namespace {

struct GridParams {
  double lo[3];
  double cs;        // cell edge (>= r * (1 + 1e-6))
  double inv_cs;    // 1 / cs: the cell index is floor((p - lo) * inv_cs) -- monotone in p like the quotient, and every
                    // kernel uses this one expression, so |dx| <= r < cs still never skips a cell
  int dim[3];
  int cells_per_graph;
};

__device__ __forceinline__ unsigned f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

__global__ void bbox_init_kernel(unsigned* bbox) {
  if (threadIdx.x < 3) bbox[threadIdx.x] = 0xFFFFFFFFu;          // min
  else if (threadIdx.x < 6) bbox[threadIdx.x] = 0u;              // max
}

__global__ void bbox_kernel(const float* __restrict__ pos, int stride, int n, unsigned* __restrict__ bbox) {
  unsigned mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      unsigned o = f2ord(pos[(size_t)i * stride + d]);
      mn[d] = min(mn[d], o);
      mx[d] = max(mx[d], o);
    }
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[d] = min(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
      mx[d] = max(mx[d], __shfl_xor_sync(0xffffffffu, mx[d], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      atomicMin(&bbox[d], mn[d]);
      atomicMax(&bbox[3 + d], mx[d]);
    }
  }
}

// One thread: choose the cell edge so that (cells per graph) * n_graphs fits the cell budget.
__device__ void grid_setup_body(const unsigned* bbox, double r, int n_graphs, int max_cells, GridParams* gp) {
  double lo[3], ext[3];
  for (int d = 0; d < 3; ++d) {
    lo[d] = (double)ord2f(bbox[d]);
    double hi = (double)ord2f(bbox[3 + d]);
    ext[d] = hi - lo[d];
    if (!(ext[d] >= 0.0)) ext[d] = 0.0;  // empty / NaN input
  }
  double cs = r * (1.0 + 1e-6);
  if (!(cs > 0.0)) cs = 1.0;
  int dim[3];
  const double budget = (double)max_cells / (double)(n_graphs > 0 ? n_graphs : 1);
  for (int it = 0; it < 200; ++it) {
    double prod = 1.0;
    for (int d = 0; d < 3; ++d) {
      double c = floor(ext[d] / cs) + 1.0;
      if (c > 2.0e6) c = 2.0e6;
      dim[d] = (int)c;
      prod *= c;
    }
    if (prod <= budget) break;
    cs *= 1.26;
  }
  for (int d = 0; d < 3; ++d) {
    gp->lo[d] = lo[d];
    gp->dim[d] = dim[d];
  }
  gp->cs = cs;
  gp->inv_cs = 1.0 / cs;
  gp->cells_per_graph = dim[0] * dim[1] * dim[2];
}

__global__ void grid_setup_kernel(const unsigned* __restrict__ bbox, double r, int n_graphs, int max_cells,
                                  GridParams* __restrict__ gp) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  grid_setup_body(bbox, r, n_graphs, max_cells, gp);
}

__device__ __forceinline__ void cell_coords(const GridParams& gp, float x, float y, float z, int c[3]) {
  const float p[3] = {x, y, z};
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    // fp64 keeps floor((p - lo) / cs) monotone, so |dx| <= r < cs never skips a cell (a multiplication by 1 / cs:
    // fp64 division is microcoded and dominated the single-block cell-list kernel)
    int v = (int)floor(((double)p[d] - gp.lo[d]) * gp.inv_cs);
    c[d] = min(max(v, 0), gp.dim[d] - 1);
  }
}

// Tiny inputs (n <= kTinyN, e.g. one 2 000-particle rollout): ONE block builds the whole cell list -- bounding box, grid
// set-up, per-cell counts and their exclusive scan in shared memory, cell-sorted positions -- and clears the in-degree /
// fill counters and the tail ticket of the neighbour kernel (replaces prepare_small + cell_count + scan + cell_fill).
constexpr int kTinyN = 8192;
constexpr int kTinyCells = 8192;
constexpr int kTinyThreads = 1024;
constexpr int kTinyPer = kTinyN / kTinyThreads;       // particles per thread

__global__ void __launch_bounds__(kTinyThreads)
cell_list_tiny_kernel(const float* __restrict__ pos, int stride, int n, int n_per_graph, double r, int n_graphs,
                      int max_cells /*capacity of cell_start - 2*/, GridParams* __restrict__ gp_out, int* __restrict__ cell_start, float4* __restrict__ sorted,
                      int* __restrict__ indeg_fill /*2(n+1)*/, int* __restrict__ tail_ticket) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  __shared__ unsigned s_box[6];
  __shared__ GridParams s_gp;
  __shared__ int s_cnt[kTinyCells];
  __shared__ int s_warp[kTinyThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 3) s_box[tid] = 0xFFFFFFFFu;
  else if (tid < 6) s_box[tid] = 0u;
  __syncthreads();
  float px[kTinyPer], py[kTinyPer], pz[kTinyPer];
  unsigned mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
#pragma unroll
  for (int k = 0; k < kTinyPer; ++k) {
    const int i = tid + k * kTinyThreads;
    px[k] = py[k] = pz[k] = 0.f;
    if (i < n) {
      const float* p = pos + (size_t)i * stride;
      px[k] = p[0]; py[k] = p[1]; pz[k] = p[2];
      const unsigned o[3] = {f2ord(px[k]), f2ord(py[k]), f2ord(pz[k])};
#pragma unroll
      for (int d = 0; d < 3; ++d) { mn[d] = min(mn[d], o[d]); mx[d] = max(mx[d], o[d]); }
    }
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[d] = min(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
      mx[d] = max(mx[d], __shfl_xor_sync(0xffffffffu, mx[d], o));
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int d = 0; d < 3; ++d) { atomicMin(&s_box[d], mn[d]); atomicMax(&s_box[3 + d], mx[d]); }
  }
  for (int i = tid; i < kTinyCells; i += kTinyThreads) s_cnt[i] = 0;
  for (int i = tid; i < 2 * (n + 1); i += kTinyThreads) indeg_fill[i] = 0;
  if (tid == 0) *tail_ticket = 0;
  __syncthreads();
  if (tid == 0) {
    grid_setup_body(s_box, r, n_graphs, min(kTinyCells, max_cells), &s_gp);   // cells <= both the shared counters and cell_start
    *gp_out = s_gp;
  }
  __syncthreads();
  const GridParams gp = s_gp;
  const int cells = gp.cells_per_graph * n_graphs;      // <= kTinyCells by construction of the grid
  int cell[kTinyPer];
#pragma unroll
  for (int k = 0; k < kTinyPer; ++k) {
    const int i = tid + k * kTinyThreads;
    cell[k] = -1;
    if (i < n) {
      int c[3];
      cell_coords(gp, px[k], py[k], pz[k], c);
      cell[k] = (i / n_per_graph) * gp.cells_per_graph + (c[2] * gp.dim[1] + c[1]) * gp.dim[0] + c[0];
      atomicAdd(&s_cnt[cell[k]], 1);
    }
  }
  __syncthreads();
  // exclusive scan of s_cnt[0 .. cells): 8 consecutive cells per thread
  {
    constexpr int kPerT = kTinyCells / kTinyThreads;
    int v[kPerT], sum = 0;
#pragma unroll
    for (int k = 0; k < kPerT; ++k) { v[k] = s_cnt[tid * kPerT + k]; sum += v[k]; }
    int x = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      s_warp[lane] = w;
    }
    __syncthreads();
    int ex = (warp == 0 ? 0 : s_warp[warp - 1]) + x - sum;
#pragma unroll
    for (int k = 0; k < kPerT; ++k) {
      const int c = tid * kPerT + k;
      s_cnt[c] = ex;                                   // becomes the fill cursor of the cell
      if (c <= cells) cell_start[c] = ex;              // cell_start[cells] == n
      ex += v[k];
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kTinyPer; ++k) {
    const int i = tid + k * kTinyThreads;
    if (i < n) sorted[atomicAdd(&s_cnt[cell[k]], 1)] = make_float4(px[k], py[k], pz[k], __int_as_float(i));
  }
}

// Small inputs (n <= kSmallN): one block computes the bounding box, sets the grid up and zeroes the counters
// (replaces bbox_init + bbox + grid_setup + three memsets).
constexpr int kSmallN = 1 << 15;
__device__ void grid_setup_body(const unsigned* bbox, double r, int n_graphs, int max_cells, GridParams* gp);

__global__ void __launch_bounds__(1024)
prepare_small_kernel(const float* __restrict__ pos, int stride, int n, double r, int n_graphs, int max_cells,
                     GridParams* __restrict__ gp, int* __restrict__ cell_cnt, int* __restrict__ indeg_fill /*2(n+1)*/) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  __shared__ unsigned s_box[6];
  if (threadIdx.x < 3) s_box[threadIdx.x] = 0xFFFFFFFFu;
  else if (threadIdx.x < 6) s_box[threadIdx.x] = 0u;
  __syncthreads();
  unsigned mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const unsigned o = f2ord(pos[(size_t)i * stride + d]);
      mn[d] = min(mn[d], o);
      mx[d] = max(mx[d], o);
    }
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[d] = min(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
      mx[d] = max(mx[d], __shfl_xor_sync(0xffffffffu, mx[d], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      atomicMin(&s_box[d], mn[d]);
      atomicMax(&s_box[3 + d], mx[d]);
    }
  }
  for (int i = threadIdx.x; i <= max_cells; i += blockDim.x) cell_cnt[i] = 0;
  for (int i = threadIdx.x; i < 2 * (n + 1); i += blockDim.x) indeg_fill[i] = 0;
  __syncthreads();
  if (threadIdx.x == 0) grid_setup_body(s_box, r, n_graphs, max_cells, gp);
}

__global__ void cell_count_kernel(const float* __restrict__ pos, int stride, int n, int n_per_graph,
                                  const GridParams* __restrict__ gpp, int* __restrict__ cell_of,
                                  int* __restrict__ cell_cnt) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const GridParams gp = *gpp;
  const float* p = pos + (size_t)i * stride;
  int c[3];
  cell_coords(gp, p[0], p[1], p[2], c);
  int g = i / n_per_graph;
  int cell = g * gp.cells_per_graph + (c[2] * gp.dim[1] + c[1]) * gp.dim[0] + c[0];
  cell_of[i] = cell;
  atomicAdd(&cell_cnt[cell], 1);
}

// sorted[slot] = (x, y, z, node id); consumes cell_cnt back to zero.
__global__ void cell_fill_kernel(const float* __restrict__ pos, int stride, int n,
                                 const int* __restrict__ cell_of, const int* __restrict__ cell_start,
                                 int* __restrict__ cell_cnt, float4* __restrict__ sorted) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int cell = cell_of[i];
  int slot = cell_start[cell] + atomicSub(&cell_cnt[cell], 1) - 1;
  const float* p = pos + (size_t)i * stride;
  sorted[slot] = make_float4(p[0], p[1], p[2], __int_as_float(i));
}

constexpr int kNbrWarps = 8;      // warps (queries) per block
constexpr int kMaxK = 64;

__device__ __forceinline__ bool key_less(double da, int ja, double db, int jb) {
  return da < db || (da == db && ja < jb);
}

// Bitonic sort of 32 * PL (d2, j) keys held PL per lane (element e = slot * 32 + lane), ascending: afterwards the key of
// rank r sits in lane r % 32, slot r / 32.  Keys are unique (j is), padding is (1e300, INT_MAX).  Exchanges at distance
// >= 32 are register-local, the others one shuffle exchange per slot: 21 stages for PL = 2 (20 of them shuffled),
// 28 for PL = 4 -- about a third of the instructions of K warp-wide arg-min rounds.
template <int PL>
__device__ __forceinline__ void warp_bitonic_sort(double (&d)[PL], int (&j)[PL], int lane) {
  constexpr int kLog = PL == 2 ? 6 : 7;           // log2(32 * PL)
#pragma unroll
  for (int lk = 1; lk <= kLog; ++lk) {
    const int k = 1 << lk;
#pragma unroll
    for (int ls = kLog - 1; ls >= 0; --ls) {
      if (ls >= lk) continue;                       // (fixed trip count so that both loops unroll completely)
      const int st = 1 << ls;
      if (st >= 32) {
        const int ss = st >> 5;
#pragma unroll
        for (int s0 = 0; s0 < PL; ++s0) {
          if ((s0 & ss) == 0) {
            const int s1 = s0 | ss;
            const bool asc = (((s0 << 5) & k) == 0);     // bit of e selected by k: slot bits only (k >= 64 here)
            const bool lt = key_less(d[s1], j[s1], d[s0], j[s0]);
            if (lt == asc) {
              const double td = d[s0]; d[s0] = d[s1]; d[s1] = td;
              const int tj = j[s0]; j[s0] = j[s1]; j[s1] = tj;
            }
          }
        }
      } else {
#pragma unroll
        for (int s0 = 0; s0 < PL; ++s0) {
          const double od = __shfl_xor_sync(0xffffffffu, d[s0], st);
          const int oj = __shfl_xor_sync(0xffffffffu, j[s0], st);
          const int e = (s0 << 5) | lane;
          const bool asc = (e & k) == 0, lower = (lane & st) == 0;
          // the lower element of an ascending pair keeps the smaller key
          const bool take = key_less(od, oj, d[s0], j[s0]) == (asc == lower);
          if (take) { d[s0] = od; j[s0] = oj; }
        }
      }
    }
  }
}

// buffer -> registers (padded with (1e300, INT_MAX)), sorted
template <int PL>
__device__ __forceinline__ void load_and_sort(const double* bd, const int* bj, int count, int lane, double (&kd)[PL],
                                              int (&kj)[PL]) {
#pragma unroll
  for (int i = 0; i < PL; ++i) {
    const int e = lane + 32 * i;
    kd[i] = e < count ? bd[e] : 1e300;
    kj[i] = e < count ? bj[e] : 0x7fffffff;
  }
  warp_bitonic_sort<PL>(kd, kj, lane);
}

// One warp per query (cell-sorted order).  Candidates within the admission bound are appended to a per-warp buffer of
// 32 * PL keys; when the next 32 candidates might not fit, the buffer is sorted, its K smallest kept and the bound
// lowered to the K-th smallest distance (a candidate farther than K known ones can never make the list; ties at
// the bound stay in and are resolved by index).  PL = 2 serves K <= 32, PL = 4 K <= 64.
template <int PL>
__global__ void __launch_bounds__(kNbrWarps * 32)
neighbour_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                 const GridParams* __restrict__ gpp, int n, int n_per_graph, double r2, int K,
                 int* __restrict__ nbr, int* __restrict__ cnt_out, int* __restrict__ indeg) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  constexpr int kBuf = 32 * PL;
  __shared__ double s_bd[kNbrWarps][kBuf];
  __shared__ int s_bj[kNbrWarps][kBuf];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int q = blockIdx.x * kNbrWarps + w;  // query in cell-sorted order (neighbouring warps share cells)
  if (q >= n) return;
  const GridParams gp = *gpp;
  double* bd = s_bd[w];
  int* bj = s_bj[w];

  const float4 me = sorted[q];
  const int qi = __float_as_int(me.w);
  const double px = (double)me.x, py = (double)me.y, pz = (double)me.z;
  int c[3];
  cell_coords(gp, me.x, me.y, me.z, c);
  const int gbase = (qi / n_per_graph) * gp.cells_per_graph;

  // The 27 neighbouring cells are 9 contiguous runs of the cell-sorted array (3 cells along x each): lane l < 9
  // fetches run l; the runs are then walked as ONE flattened candidate list, 32 candidates per step, with the
  // next step's positions already in flight while the current ones are tested.
  int beg = 0, len = 0;
  if (lane < 9) {
    const int z = c[2] + lane / 3 - 1, y = c[1] + lane % 3 - 1;
    if (z >= 0 && z < gp.dim[2] && y >= 0 && y < gp.dim[1]) {
      const int x0 = max(c[0] - 1, 0), x1 = min(c[0] + 1, gp.dim[0] - 1);
      const int row = gbase + (z * gp.dim[1] + y) * gp.dim[0];
      beg = cell_start[row + x0];
      len = cell_start[row + x1 + 1] - beg;
    }
  }
  int incl = len;
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 8);
  int rb[9], rx[9];                               // run start in `sorted`, exclusive prefix of the run lengths
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    rb[k] = __shfl_sync(0xffffffffu, beg, k);
    rx[k] = __shfl_sync(0xffffffffu, incl - len, k);
  }
  auto fetch = [&](int f) -> float4 {
    if (f >= total) return make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
    int sb = rb[0], sx = rx[0];
#pragma unroll
    for (int t = 1; t < 9; ++t)
      if (f >= rx[t]) { sb = rb[t]; sx = rx[t]; }   // the last run whose prefix is <= f (empty runs are passed over)
    return sorted[sb + (f - sx)];
  };
  double kd[PL];
  int kj[PL];

  int count = 0;
  double bound = r2;
  float4 nxt = fetch(lane);
  for (int f0 = 0; f0 < total; f0 += 32) {
    const float4 o = nxt;
    nxt = fetch(f0 + 32 + lane);
    const bool valid = f0 + lane < total;
    const double dx = px - (double)o.x, dyy = py - (double)o.y, dzz = pz - (double)o.z;
    const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dyy, dyy)), __dmul_rn(dzz, dzz));
    const bool in = valid && d2 <= bound;
    const unsigned m = __ballot_sync(0xffffffffu, in);
    if (in) {
      const int at = count + __popc(m & ((1u << lane) - 1u));
      bd[at] = d2;
      bj[at] = __float_as_int(o.w);
    }
    count += __popc(m);
    __syncwarp();
    if (count + 32 > kBuf && f0 + 32 < total) {  // keep only the K best so far, then continue scanning
      load_and_sort<PL>(bd, bj, count, lane, kd, kj);
      const int kept = count < K ? count : K;
      __syncwarp();
#pragma unroll
      for (int i = 0; i < PL; ++i) {
        const int e = lane + 32 * i;
        if (e < kept) { bd[e] = kd[i]; bj[e] = kj[i]; }
      }
      count = kept;
      if (kept == K) {
        // (every slot is shuffled and the right one selected afterwards: indexing kd[] with (K - 1) >> 5 would put the
        // key registers into local memory)
        double kth = __shfl_sync(0xffffffffu, kd[0], (K - 1) & 31);
#pragma unroll
        for (int i = 1; i < PL; ++i) {
          const double t = __shfl_sync(0xffffffffu, kd[i], (K - 1) & 31);
          kth = ((K - 1) >> 5) == i ? t : kth;
        }
        bound = kth;
      }
      __syncwarp();
    }
  }
  load_and_sort<PL>(bd, bj, count, lane, kd, kj);
  const int kept = count < K ? count : K;
#pragma unroll
  for (int i = 0; i < PL; ++i) {
    const int e = lane + 32 * i;
    if (e < K) {
      const int j = e < kept ? kj[i] : -1;
      nbr[(size_t)qi * K + e] = j;
      if (j >= 0) atomicAdd(&indeg[j], 1);
    }
  }
  if (lane == 0) cnt_out[qi] = kept;
}

// K <= 32 fits two keys per lane (buffer 64 >= K + 32), larger K four
static cudaError_t launch_neighbour(cudaStream_t stream, const float4* sorted, const int* cell_start, const GridParams* gp,
                                    int n, int n_per_graph, double r2, int K, int* nbr, int* cnt, int* indeg) {
  if (K <= 32)
    return launch_pdl(neighbour_kernel<2>, ceil_div(n, kNbrWarps), kNbrWarps * 32, 0, stream, 1, sorted, cell_start, gp, n,
                      n_per_graph, r2, K, nbr, cnt, indeg);
  return launch_pdl(neighbour_kernel<4>, ceil_div(n, kNbrWarps), kNbrWarps * 32, 0, stream, 1, sorted, cell_start, gp, n,
                    n_per_graph, r2, K, nbr, cnt, indeg);
}

// ---- receiver CSR over "edge slots" (slot -> receiver, -1 = empty slot) ----
__global__ void indeg_count_kernel(const int* __restrict__ recv_of_slot, long long n_slots,
                                   const int* __restrict__ n_valid_dev, int* __restrict__ indeg) {
  long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slots) return;
  if (n_valid_dev && s >= *n_valid_dev) return;
  int j = recv_of_slot[s];
  if (j >= 0) atomicAdd(&indeg[j], 1);
}

__global__ void recv_fill_kernel(const int* __restrict__ recv_of_slot, long long n_slots,
                                 const int* __restrict__ n_valid_dev, const int* __restrict__ recv_ptr,
                                 int* __restrict__ fill, int* __restrict__ keys) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slots) return;
  if (n_valid_dev && s >= *n_valid_dev) return;
  int j = recv_of_slot[s];
  if (j < 0) return;
  int at = recv_ptr[j] + atomicAdd(&fill[j], 1);
  keys[at] = (int)s;
}

// Warp per receiver: every lane ranks its one or two keys against the whole segment with shuffles and
// writes the sorted edge straight out (keys are distinct slot ids, so ranks are a permutation).
__global__ void __launch_bounds__(256)
recv_sort_emit_warp_kernel(const int* __restrict__ recv_ptr, int n, const int* __restrict__ keys, int K,
                           const int* __restrict__ edge_ptr, const int* __restrict__ send_of_slot,
                           int* __restrict__ esend, int* __restrict__ erecv, int* __restrict__ eref) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (j >= n) return;
  const int b = recv_ptr[j], len = recv_ptr[j + 1] - b;
  const int k0 = lane < len ? keys[b + lane] : 0x7FFFFFFF;
  const int k1 = lane + 32 < len ? keys[b + lane + 32] : 0x7FFFFFFF;
  int r0 = 0, r1 = 0;
#pragma unroll 8
  for (int o = 0; o < 32; ++o) {
    const int a0 = __shfl_sync(0xffffffffu, k0, o);
    r0 += a0 < k0; r1 += a0 < k1;
  }
  if (len > 32) {
#pragma unroll 8
    for (int o = 0; o < 32; ++o) {
      const int a1 = __shfl_sync(0xffffffffu, k1, o);
      r0 += a1 < k0; r1 += a1 < k1;
    }
  }
  auto emit = [&](int slot, int rank) {
    int snd, ref;
    if (K > 0) {
      snd = slot / K;
      ref = edge_ptr[snd] + (slot - snd * K);
    } else {
      snd = send_of_slot[slot];
      ref = slot;
    }
    esend[b + rank] = snd;
    erecv[b + rank] = j;
    eref[b + rank] = ref;
  };
  if (len <= 64) {
    if (lane < len) emit(k0, r0);
    if (lane + 32 < len) emit(k1, r1);
  } else {                                   // rare (the cap bounds out-degree, not in-degree): rank against global memory
    for (int t = lane; t < len; t += 32) {
      const int k = keys[b + t];
      int rank = 0;
      for (int u = 0; u < len; ++u) rank += keys[b + u] < k;
      emit(k, rank);
    }
  }
}

// tiny inputs: out-degrees -> edge_ptr (+ n_edges) and in-degrees -> recv_ptr in ONE launch of two blocks
__global__ void __launch_bounds__(kScanThreads)
scan_degrees_kernel(const int* __restrict__ cnt, const int* __restrict__ indeg, int n, int* __restrict__ edge_ptr,
                    int* __restrict__ n_edges, int* __restrict__ recv_ptr) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  __shared__ int s_scan[9];
  const int* in = blockIdx.x == 0 ? cnt : indeg;
  int* out = blockIdx.x == 0 ? edge_ptr : recv_ptr;
  int carry = 0;
  for (int base0 = 0; base0 < n; base0 += kScanTile) {
    const int base = base0 + (int)threadIdx.x * kScanItems;
    int v[kScanItems], sum = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
      v[i] = base + i < n ? in[base + i] : 0;
      sum += v[i];
    }
    int total;
    int ex = block_exclusive_scan_256(sum, &total, s_scan) + carry;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
      if (base + i < n) out[base + i] = ex;
      ex += v[i];
    }
    carry += total;
  }
  if (threadIdx.x == 0) {
    out[n] = carry;
    if (blockIdx.x == 0) *n_edges = carry;
  }
}

__global__ void copy_last_kernel(const int* __restrict__ ptr, int n, int* __restrict__ out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = ptr[n];
}

}  // namespace

// ------------------------------------------------------------------------------------------------
size_t radius_graph_workspace_bytes(int n_total, int n_graphs, int K) {
  Arena a(nullptr, 0);
  RadiusGraphWs ws;
  radius_graph_carve(a, n_total, n_graphs, K, ws);
  return a.off + 256;
}

int radius_graph_max_cells(int n_total) {
  long long c = 4LL * n_total;
  if (c < 4096) c = 4096;
  if (c > (1LL << 22)) c = 1LL << 22;
  return (int)c;
}

void radius_graph_carve(Arena& a, int n_total, int /*n_graphs*/, int K, RadiusGraphWs& ws) {
  ws.max_cells = radius_graph_max_cells(n_total);
  ws.bbox = a.take<unsigned>(8);
  ws.grid = a.take<char>(sizeof(GridParams));
  ws.cell_of = a.take<int>(n_total);
  ws.cell_cnt = a.take<int>((size_t)ws.max_cells + 1);
  ws.cell_start = a.take<int>((size_t)ws.max_cells + 2);
  ws.sorted = a.take<float4>(n_total);
  ws.indeg = a.take<int>(2 * ((size_t)n_total + 1));   // indeg and fill are contiguous: zeroed together
  ws.fill = ws.indeg + ((size_t)n_total + 1);
  ws.keys = a.take<int>((size_t)n_total * K);
  size_t scan_n = (size_t)ws.max_cells + 1 > (size_t)n_total + 1 ? (size_t)ws.max_cells + 1 : (size_t)n_total + 1;
  ws.scan_ws = a.take<int>(scan_workspace_ints((long long)scan_n) + 8);
}

int build_recv_csr(const int* recv_of_slot, long long n_slots, const int* n_valid_dev, int n_nodes, int K,
                   const int* edge_ptr, const int* send_of_slot, int* indeg, int* fill, int* keys, int* scan_ws,
                   bool indeg_ready, int* recv_ptr, int* esend, int* erecv, int* eref, cudaStream_t stream,
                   bool fill_ready, bool recv_ptr_ready) {
  const int T = 256;
  if (!indeg_ready) {
    GM_CUDA(cudaMemsetAsync(indeg, 0, sizeof(int) * ((size_t)n_nodes + 1), stream));
    if (n_slots > 0) {
      indeg_count_kernel<<<ceil_div(n_slots, T), T, 0, stream>>>(recv_of_slot, n_slots, n_valid_dev, indeg);
      GM_LAUNCHED(1);
    }
  }
  if (!fill_ready) GM_CUDA(cudaMemsetAsync(fill, 0, sizeof(int) * ((size_t)n_nodes + 1), stream));
  if (!recv_ptr_ready) {
    int rc = exclusive_scan_i32(indeg, recv_ptr, n_nodes, scan_ws, stream);
    if (rc) return rc;
  }
  if (n_slots > 0) {
    GM_CUDA(launch_pdl(recv_fill_kernel, (int)ceil_div(n_slots, T), T, 0, stream, 1, recv_of_slot, n_slots, n_valid_dev, recv_ptr,
                       fill, keys));
    GM_LAUNCHED(1);
  }
  if (n_nodes > 0) {
    GM_CUDA(launch_pdl(recv_sort_emit_warp_kernel, ceil_div(n_nodes, 8), 256, 0, stream, 1, recv_ptr, n_nodes, keys, K, edge_ptr,
                       send_of_slot, esend, erecv, eref));
    GM_LAUNCHED(1);
  }
  GM_LAUNCH_CHECK("build_recv_csr");
  return GM_OK;
}

int radius_graph(const float* pos, int pos_stride, int n_per_graph, int n_graphs, double r, int K, int* nbr,
                 int* cnt, int* edge_ptr, int* recv_ptr, int* esend, int* erecv, int* eref, int* n_edges,
                 const RadiusGraphWs& ws, cudaStream_t stream) {
  const long long n_ll = (long long)n_per_graph * n_graphs;
  GM_REQUIRE(n_per_graph > 0 && n_graphs > 0, "radius_graph: empty input (n_per_graph=%d n_graphs=%d)",
             n_per_graph, n_graphs);
  GM_REQUIRE(K >= 1 && K <= kMaxK, "radius_graph: max_neighbours=%d outside [1,%d]", K, kMaxK);
  GM_REQUIRE(n_ll * K < (1LL << 31), "radius_graph: n*K=%lld overflows int32 slots", n_ll * K);
  GM_REQUIRE(r > 0.0, "radius_graph: conn_r must be positive");
  const int n = (int)n_ll;
  const int T = 256;
  GridParams* gp = reinterpret_cast<GridParams*>(ws.grid);

  const bool small = n <= kSmallN;
  static const bool tiny_off = [] { const char* e = getenv("GM_GRAPH_TINY"); return e && e[0] == '0'; }();   // A/B aid
  const bool tiny = n <= kTinyN && kNbrWarps * 32 == kScanThreads && !tiny_off;
  GM_PROF(kProfGrid);
  if (tiny) {
    // one block builds the cell list, one launch scans both degree arrays: 5 launches for the whole graph
    int* ticket = reinterpret_cast<int*>(ws.bbox) + 7;
    GM_CUDA(launch_pdl(cell_list_tiny_kernel, 1, kTinyThreads, 0, stream, 1, pos, pos_stride, n, n_per_graph, r, n_graphs,
                       ws.max_cells, gp,
                       ws.cell_start, ws.sorted, ws.indeg, ticket));
    GM_PROF(kProfNeighbour);
    GM_CUDA(launch_neighbour(stream, ws.sorted, ws.cell_start, gp, n, n_per_graph, r * r, K, nbr, cnt, ws.indeg));
    GM_PROF(kProfCsr);
    GM_CUDA(launch_pdl(scan_degrees_kernel, 2, kScanThreads, 0, stream, 1, (const int*)cnt, (const int*)ws.indeg, n, edge_ptr,
                       n_edges, recv_ptr));
    GM_LAUNCHED(3);
    GM_LAUNCH_CHECK("radius_graph kernels (tiny path)");
    return build_recv_csr(nbr, (long long)n * K, nullptr, n, K, edge_ptr, nullptr, ws.indeg, ws.fill, ws.keys, ws.scan_ws,
                          /*indeg_ready=*/true, recv_ptr, esend, erecv, eref, stream, /*fill_ready=*/true,
                          /*recv_ptr_ready=*/true);
  }
  if (small) {
    GM_CUDA(launch_pdl(prepare_small_kernel, 1, 1024, 0, stream, 1, pos, pos_stride, n, r, n_graphs, ws.max_cells, gp, ws.cell_cnt,
                       ws.indeg));
    GM_LAUNCHED(1);
  } else {
    bbox_init_kernel<<<1, 32, 0, stream>>>(ws.bbox);
    bbox_kernel<<<min(ceil_div(n, T), 4 * kNumSMs), T, 0, stream>>>(pos, pos_stride, n, ws.bbox);
    grid_setup_kernel<<<1, 32, 0, stream>>>(ws.bbox, r, n_graphs, ws.max_cells, gp);
    GM_CUDA(cudaMemsetAsync(ws.cell_cnt, 0, sizeof(int) * ((size_t)ws.max_cells + 1), stream));
    GM_CUDA(cudaMemsetAsync(ws.indeg, 0, sizeof(int) * ((size_t)n + 1), stream));
    GM_LAUNCHED(3);
  }
  GM_CUDA(launch_pdl(cell_count_kernel, ceil_div(n, T), T, 0, stream, 1, pos, pos_stride, n, n_per_graph, gp, ws.cell_of,
                     ws.cell_cnt));
  // only the cells of the live grid are scanned (device-side bound; the workspace is sized for the worst case)
  int rc = exclusive_scan_i32(ws.cell_cnt, ws.cell_start, ws.max_cells, ws.scan_ws, stream, nullptr,
                              &gp->cells_per_graph, n_graphs);
  if (rc) return rc;
  GM_CUDA(launch_pdl(cell_fill_kernel, ceil_div(n, T), T, 0, stream, 1, pos, pos_stride, n, ws.cell_of, ws.cell_start,
                     ws.cell_cnt, ws.sorted));
  GM_PROF(kProfNeighbour);
  GM_CUDA(launch_neighbour(stream, ws.sorted, ws.cell_start, gp, n, n_per_graph, r * r, K, nbr, cnt, ws.indeg));
  GM_LAUNCHED(3);
  GM_LAUNCH_CHECK("radius_graph kernels");
  GM_PROF(kProfCsr);
  rc = exclusive_scan_i32(cnt, edge_ptr, n, ws.scan_ws, stream, n_edges);
  if (rc) return rc;
  return build_recv_csr(nbr, (long long)n * K, nullptr, n, K, edge_ptr, nullptr, ws.indeg, ws.fill, ws.keys,
                        ws.scan_ws, /*indeg_ready=*/true, recv_ptr, esend, erecv, eref, stream, /*fill_ready=*/small, false);
}

GM_TRACE_DUMP(trace_dump_radius)

}  // namespace gm
