// Shared device helpers for the sm_100a kernels: mbarrier, TMA, tcgen05/TMEM,
// programmatic dependent launch, vector loads, warp reductions.
// Everything here is inline PTX written for sm_100a only (no fallback paths).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace bcg {

// ----------------------------------------------------------------------------
// status codes returned by every extern "C" entry point (include/bcg_b200.h)
// ----------------------------------------------------------------------------
enum : int {
  BCG_OK = 0,
  BCG_ERR_ARG = 1,       // bad argument (shape / alignment / null)
  BCG_ERR_CUDA = 2,      // a CUDA runtime / driver call failed
  BCG_ERR_UNSUPPORTED = 3,
  BCG_ERR_TOO_LARGE = 4, // exceeds a compiled-in maximum (BCG_STATUS_TOO_LARGE in include/bcg_b200.h)
};

#define BCG_CUDA_CHECK(expr)                                   \
  do {                                                         \
    cudaError_t _e = (expr);                                   \
    if (_e != cudaSuccess) {                                   \
      bcg::set_last_error(#expr, cudaGetErrorString(_e));      \
      return bcg::BCG_ERR_CUDA;                                \
    }                                                          \
  } while (0)

void set_last_error(const char* what, const char* detail);

// Launch flags of the calling thread's current composite step (set by model_decode): when
// BCG_GEMM_PDL is on, every kernel of the step is launched with programmatic stream
// serialisation, i.e. it may start while its predecessor is still running; each kernel
// announces its dependents at entry (griddepcontrol.launch_dependents) and waits for its
// predecessor's results with griddepcontrol.wait before touching them.
int& launch_flags();

// An SM runs CTAs of different kernels side by side only if they agree on its L1 / shared-memory
// split.  The GEMM and attention CTAs need the largest shared-memory carve-out; a small kernel left
// at its default (mostly-L1) split cannot become resident next to them, so under programmatic
// dependent launch it would start only after they exit — and so would ITS dependents.  Every
// kernel launched through launch_k therefore asks for the same maximum carve-out (once).
void prefer_max_smem_carveout(const void* kernel);
// Raises the kernel's dynamic shared-memory limit on the CURRENT device (once per kernel and device).
int ensure_dynamic_smem(const void* kernel, int bytes);

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                            cudaStream_t stream, Args&&... args) {
  prefer_max_smem_carveout(reinterpret_cast<const void*>(kernel));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = (launch_flags() & 1) ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ----------------------------------------------------------------------------
// shared-memory address / mbarrier
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_acq_rel_gpu() {
  asm volatile("fence.acq_rel.gpu;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded spin: a protocol bug traps (reported as a CUDA error to the caller)
// instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) __trap();
  }
}

// ----------------------------------------------------------------------------
// TMA (cp.async.bulk.tensor) — 2D tiled loads into 128B-swizzled smem
// ----------------------------------------------------------------------------
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;

__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, uint64_t* bar,
                                            int c0, int c1, uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
// 1D bulk copy global -> shared (UBLKCP), completes on an mbarrier.
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// ----------------------------------------------------------------------------
// tcgen05 / TMEM
// ----------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 inputs, fp32 accumulate; one thread issues.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives once all previously issued MMAs of this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 32 lanes x 32 bit, 16 consecutive columns: thread t of the warp gets lane (base_lane + t).
__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

// K-major, 128B-swizzled operand tile (rows x 64 bf16, 8-row atoms of 1024 B):
// start>>4 | LBO=1 | SBO=1024>>4 | version=1 (bit 46) | SWIZZLE_128B (2 << 61).
// (field layout: cute::UMMA::SmemDescriptor, CUTLASS cute/arch/mma_sm100_desc.hpp)
__device__ __forceinline__ uint64_t make_kmajor_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
// kind::f16 instruction descriptor: D=f32, A=B=bf16, both K-major, M=128, N=n.
__host__ __device__ constexpr uint32_t make_idesc_bf16_m128(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(n >> 3) << 17) |
         (static_cast<uint32_t>(128 >> 4) << 24);
}

// ----------------------------------------------------------------------------
// ----------------------------------------------------------------------------
// Timeline instrumentation (dev tool, tools/timeline.py): compiled in only with -DBCG_TIMELINE,
// into a separate library.  Chosen threads stamp %globaltimer into a per-translation-unit ring;
// BCG_TL_DEFINE_DUMP(name) exports that unit's dump/reset function.
// ----------------------------------------------------------------------------
#ifdef BCG_TIMELINE
constexpr unsigned kTlCap = 1u << 18;
static __device__ unsigned long long g_tl[kTlCap];
static __device__ unsigned int g_tl_n;
__device__ __forceinline__ void tl_mark(unsigned id, unsigned phase) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  const unsigned blk = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  const unsigned i = atomicAdd(&g_tl_n, 1u);
  if (i < kTlCap) g_tl[i] = (t << 20) | (static_cast<unsigned long long>(id) << 16) |
                            (static_cast<unsigned long long>(phase) << 12) | (blk < 4095u ? blk : 4095u);
}
#define BCG_TL(id, phase) ::bcg::tl_mark(id, phase)
#define BCG_TL_DEFINE_DUMP(name)                                                              \
  extern "C" int name(unsigned long long* host, int cap, int reset) {                         \
    unsigned n = 0;                                                                            \
    if (cudaMemcpyFromSymbol(&n, ::bcg::g_tl_n, sizeof(n)) != cudaSuccess) return -1;          \
    if (n > ::bcg::kTlCap) n = ::bcg::kTlCap;                                                  \
    if (static_cast<int>(n) > cap) n = cap;                                                    \
    if (n && cudaMemcpyFromSymbol(host, ::bcg::g_tl, n * sizeof(unsigned long long)) != cudaSuccess) return -1; \
    if (reset) {                                                                               \
      const unsigned z = 0;                                                                    \
      if (cudaMemcpyToSymbol(::bcg::g_tl_n, &z, sizeof(z)) != cudaSuccess) return -1;          \
    }                                                                                          \
    return static_cast<int>(n);                                                                \
  }
#else
#define BCG_TL(id, phase) ((void)0)
#define BCG_TL_DEFINE_DUMP(name)
#endif

// programmatic dependent launch (PDL)
// ----------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ----------------------------------------------------------------------------
// misc
// ----------------------------------------------------------------------------
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "selp.b32 %0, 1, 0, P;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ float bf16lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

}  // namespace bcg
