// Shared device/host helpers for the gnn-manip B200 hot path (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "gnn_manip_b200 kernels are written for sm_100a (B200) only"
#endif

namespace gm {

// ---------------------------------------------------------------------------------------------
// status / error reporting across the C ABI (no exceptions cross it)
// ---------------------------------------------------------------------------------------------
enum Status : int {
  GM_OK = 0,
  GM_ERR_INVALID = 1,     // bad argument
  GM_ERR_WORKSPACE = 2,   // workspace too small
  GM_ERR_CUDA = 3,        // CUDA runtime error (see gm_last_error)
  GM_ERR_UNSUPPORTED = 4  // configuration outside the hand-written kernels
};

void set_error(const char* fmt, ...);
// number of kernels of this library launched by the calling thread (bench.py's gpu_launches)
extern thread_local long long g_launch_count;
int cuda_fail(cudaError_t e, const char* what);

#define GM_CUDA(call)                                        \
  do {                                                       \
    cudaError_t _e = (call);                                 \
    if (_e != cudaSuccess) return gm::cuda_fail(_e, #call);  \
  } while (0)

#define GM_LAUNCH_CHECK(name)                                   \
  do {                                                          \
    cudaError_t _e = cudaGetLastError();                        \
    if (_e != cudaSuccess) return gm::cuda_fail(_e, name);      \
  } while (0)
// every kernel launch site bumps the per-thread launch counter
#define GM_LAUNCHED(n) (gm::g_launch_count += (n))

#define GM_REQUIRE(cond, ...)            \
  do {                                   \
    if (!(cond)) {                       \
      gm::set_error(__VA_ARGS__);        \
      return gm::GM_ERR_INVALID;         \
    }                                    \
  } while (0)

constexpr int kNumSMs = 148;

// Per-phase CUDA-event timing of one un-captured rollout step / forward (gm_rollout_step_profile,
// gm_model_forward_profile): while a Profiler is installed on the calling thread, every launcher marks the phase it is
// about to enqueue; the time between two marks is attributed to the first one's phase.
enum ProfPhase : int {
  kProfControl = 0, kProfNodeFeat, kProfGrid, kProfNeighbour, kProfCsr, kProfEncEdge, kProfEncNode, kProfMpEdge,
  kProfMpNode, kProfIntegrate, kProfCount
};
struct Profiler {
  cudaStream_t st = nullptr;
  int n = 0;
  static constexpr int kMax = 96;
  cudaEvent_t ev[kMax];
  int cls[kMax];
  int mark(int c);             // records an event on st; c < 0 = end marker
  int finish(float* ms_by_phase, int* launches_by_phase);   // synchronises st, accumulates, destroys the events
};
extern thread_local Profiler* g_profiler;
#define GM_PROF(c) do { if (gm::g_profiler) { int _rc = gm::g_profiler->mark(c); if (_rc) return _rc; } } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-device attribute: a "configured" flag per device (the
// flags are atomics: two host threads may race to configure, which is harmless -- the call is idempotent).
struct PerDeviceOnce {
  static constexpr int kMaxDevices = 64;
  volatile int done[kMaxDevices] = {};
  // returns the current device index if it still needs configuring, -1 if already done, -2 on error
  int begin() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return -2;
    return done[dev] ? -1 : dev;
  }
  void finish(int dev) { done[dev] = 1; }
};

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// Bump allocator over a caller-provided workspace (the ABI never allocates device memory).
struct Arena {
  char* base;
  size_t size, off;
  bool ok;
  Arena(void* p, size_t n) : base((char*)p), size(n), off(0), ok(true) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 256);
    size_t bytes = count * sizeof(T);
    if (base != nullptr && off + bytes > size) ok = false;
    T* r = base ? (T*)(base + off) : nullptr;
    off += bytes;
    return r;
  }
};

#ifdef __CUDACC__
// Launch with programmatic stream serialization (and an optional cluster width): the kernel may start while its
// predecessor in the stream drains; it calls pdl_wait() before touching upstream data.
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_ex(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t stream,
                                    int cluster_x, bool programmatic, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (programmatic) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster_x;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t stream,
                                     int cluster_x, Args... args) {
  return launch_ex(kernel, grid, block, smem, stream, cluster_x, true, args...);
}

// ---------------------------------------------------------------------------------------------
// PTX wrappers: mbarrier, bulk async copy (TMA unit), proxy fences, tcgen05 / TMEM
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// mbarrier.try_wait parks the thread in hardware until the phase completes or a short system time limit expires
// (no suspend-time hint: with a hint ptxas emits NANOSLEEP back-off loops whose wake-up latency is microseconds).
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must trap (kills the context with an error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
#pragma unroll 1
  for (uint32_t spins = 0; spins < (1u << 26); ++spins)
    if (mbar_try_wait(bar, parity)) return;
  printf("gnn_manip_b200: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
  __trap();
}

// generic-proxy writes to shared memory -> visible to the async proxy (UMMA / bulk copy)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// 1-D bulk copy global -> shared through the TMA unit, completion on an mbarrier (bytes % 16 == 0).
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}

// 16-byte asynchronous global -> shared copy (LDGSTS, L2 only); src_bytes = 0 zero-fills the destination.
__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// named barrier over `nthreads` threads (ids 1..15; 0 is __syncthreads)
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- TMEM ----
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// 32 lanes x 32 columns of fp32: thread i of the warp receives TMEM lane (base_lane + i), columns [col, col+32)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

// ---- UMMA (tcgen05.mma) ----
// Shared-memory matrix descriptor, K-major operand, SWIZZLE_128B, 16-bit elements:
// rows of 128 B (64 elements), 8-row swizzle atoms of 1024 B (SBO = 1024), version 1 (sm_100).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);   // start address, bits [0,14)
  d |= (uint64_t)1 << 16;                    // leading byte offset (ignored for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;          // stride byte offset between 8-row groups
  d |= (uint64_t)1 << 46;                    // descriptor version = 1
  d |= (uint64_t)2 << 61;                    // layout type = SWIZZLE_128B
  return d;
}
// Descriptor of the operand `bytes` further on (bytes % 16 == 0): the start address sits in the low 14 bits in units of
// 16 bytes and shared memory ends below 256 KB, so this is ONE 32-bit add on the low word -- the MMA-issuing thread's
// instruction stream (not the tensor pipe) paces a chain of small MMAs, every instruction saved there shortens a GEMM.
__device__ __forceinline__ uint64_t umma_desc_advance(uint64_t desc, uint32_t bytes) {
  return desc + (uint64_t)(bytes >> 4);
}
// Instruction descriptor: kind::f16, A = B = BF16, D = F32, both K-major, M x N tile.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :
      : "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier when all previously issued MMAs of this thread have completed.
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}

// ---- programmatic dependent launch: the kernels of one model step are chained so that the next kernel's
// prologue (barrier init, TMEM allocation, parameter / weight staging -- constant data only) overlaps the tail of
// the previous kernel.  pdl_wait() must precede the first access to anything an earlier kernel of the stream wrote.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#ifdef GM_TRACE
// Debug aid (-DGM_TRACE): thread 0 of CTA 0 of every kernel records %globaltimer right after its dependency wait,
// i.e. when its predecessor completed; one buffer per translation unit, read (and reset) by gm_debug_stamps().
constexpr int kTraceCap = 160;
static __device__ unsigned long long g_trace_t[kTraceCap];
static __device__ unsigned int g_trace_n;
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    const unsigned i = atomicAdd(&g_trace_n, 1u);
    if (i < kTraceCap) g_trace_t[i] = t;
  }
}
#define GM_TRACE_DUMP(fn)                                                        \
  int fn(unsigned long long* out, int cap) {                                     \
    unsigned n = 0, z = 0;                                                       \
    cudaMemcpyFromSymbol(&n, g_trace_n, sizeof(n));                              \
    if (n > (unsigned)kTraceCap) n = kTraceCap;                                  \
    if ((int)n > cap) n = (unsigned)cap;                                         \
    if (n) cudaMemcpyFromSymbol(out, g_trace_t, sizeof(unsigned long long) * n); \
    cudaMemcpyToSymbol(g_trace_n, &z, sizeof(z));                                \
    return (int)n;                                                               \
  }
#else
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#define GM_TRACE_DUMP(fn)
#endif

// ---- bf16 hi/lo split of an fp32 pair: x ~= hi + lo with |err| <~ 2^-17 |x| ----
__device__ __forceinline__ void split_bf16x2(float a, float b, uint32_t& hi, uint32_t& lo) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  hi = *reinterpret_cast<uint32_t*>(&h);
  float ha = __uint_as_float(hi << 16);
  float hb = __uint_as_float(hi & 0xFFFF0000u);
  __nv_bfloat162 l = __floats2bfloat162_rn(a - ha, b - hb);
  lo = *reinterpret_cast<uint32_t*>(&l);
}
#endif  // __CUDACC__

}  // namespace gm
