// hetseq_b200 -- device-side building blocks for sm_100a kernels.
//
// Thin inline-PTX wrappers (mbarrier, TMA, tcgen05/TMEM), dtype helpers, Philox
// counter RNG and warp/block reductions shared by every .cu file in csrc/.
// Nothing here has a counterpart in the reference (it has no native GPU code,
// SURVEY.md section 2.4); these are the primitives the B200 kernels are built from.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

namespace hs {

// ----------------------------------------------------------------------------------------------
// error plumbing (host)
// ----------------------------------------------------------------------------------------------
void set_last_error(const char* fmt, ...);
int  check_cuda(cudaError_t e, const char* what);

#define HS_CUDA_OK(expr)                                   \
  do {                                                     \
    int _hs_rc = ::hs::check_cuda((expr), #expr);          \
    if (_hs_rc != 0) return _hs_rc;                        \
  } while (0)

// ----------------------------------------------------------------------------------------------
// small device helpers
// ----------------------------------------------------------------------------------------------
// Programmatic dependent launch (launch attribute programmaticStreamSerialization, see host.h::launch_k):
// a kernel launched that way may become resident while its predecessor in the stream is still running.
// Everything before pdl_wait() (barrier init, TMEM allocation, descriptor prefetch) overlaps the predecessor's
// tail; pdl_wait() returns once the predecessor has completed and its memory is visible, so NO global memory
// may be touched before it. pdl_launch_dependents() then lets the NEXT kernel do the same to us. Both are
// no-ops for a kernel launched without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() {
  pdl_wait();
  pdl_launch_dependents();
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one_sync() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t"
      ".reg .b32 %%rx;\n\t"
      ".reg .pred %%px;\n\t"
      "elect.sync %%rx|%%px, %1;\n\t"
      "@%%px mov.s32 %0, 1;\n\t"
      "}\n"
      : "+r"(pred)
      : "r"(0xffffffffu));
  return pred != 0;
}

// ----------------------------------------------------------------------------------------------
// mbarrier
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must trap (the launch then fails with an error) rather than hang
// the GPU. ~4e9 cycles is seconds, far beyond any legitimate wait inside one kernel.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {
      printf("hetseq_b200: mbarrier wait timed out (block %d thread %d parity %u)\n", (int)blockIdx.x,
             (int)threadIdx.x, parity);
      __trap();
    }
  }
}

// ----------------------------------------------------------------------------------------------
// TMA (cp.async.bulk.tensor) -- tiles land in shared memory and complete_tx an mbarrier
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tmap, uint64_t* bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* tmap, uint64_t* bar, int c0,
                                            int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
      "r"(c2)
      : "memory");
}

// Untiled bulk copy global -> shared (one contiguous run, 16-byte aligned, size % 16 == 0), completing on
// an mbarrier like a tensor-map load: rows of an activation matrix straight into a per-warp ring.
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// ----------------------------------------------------------------------------------------------
// tcgen05 / TMEM
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; issued by ONE thread for the whole CTA.
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- cta_group::2 (CTA pair) variants --------------------------------------------------------------
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // clears the CTA-rank bit of a shared::cluster address -> CTA 0 of the pair
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// issued by the leader CTA only; operands are read from BOTH CTAs' shared memory, D spans both TMEMs
__device__ __forceinline__ void umma_f16_2cta(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this offset in every CTA of `mask` once the issued MMAs have retired
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(mask)
      : "memory");
}
// TMA load whose completion bytes are credited to the LEADER CTA's barrier at the same offset
__device__ __forceinline__ void tma_load_2d_2cta(void* smem_dst, const CUtensorMap* tmap, uint64_t* bar, int c0,
                                                 int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0),
      "r"(c1)
      : "memory");
}
// plain arrive on the leader CTA's barrier at this offset (used by both CTAs' epilogue warps)
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerBitMask) : "memory");
}

// 32 lanes x 32 consecutive fp32 columns: thread t of the warp receives row (lane base + t).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]),
        "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
        "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// UMMA shared-memory matrix descriptor, 128-byte swizzle (sm_100 format, version field = 1).
//   bits [0,14)  start address >> 4        bits [16,30) leading byte offset >> 4
//   bits [32,46) stride byte offset >> 4   bits [46,48) version = 1     bits [61,64) layout (2 = SW128)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;
  return d;
}

// Instruction descriptor for tcgen05.mma kind::f16 with fp32 accumulation.
//   ab_fmt: 0 = fp16, 1 = bf16.  a_mn / b_mn: 1 when the operand's M/N dimension is the contiguous one.
__host__ __device__ inline uint32_t umma_idesc_f16(int ab_fmt, int a_mn, int b_mn, int m, int n) {
  uint32_t d = 0;
  d |= 1u << 4;                         // accumulator format = F32
  d |= static_cast<uint32_t>(ab_fmt) << 7;
  d |= static_cast<uint32_t>(ab_fmt) << 10;
  d |= static_cast<uint32_t>(a_mn) << 15;
  d |= static_cast<uint32_t>(b_mn) << 16;
  d |= static_cast<uint32_t>(n >> 3) << 17;
  d |= static_cast<uint32_t>(m >> 4) << 24;
  return d;
}

// ----------------------------------------------------------------------------------------------
// dtype helpers
// ----------------------------------------------------------------------------------------------
template <typename T> struct Cvt;
template <> struct Cvt<__half> {
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
  static __device__ __forceinline__ uint32_t pack2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  static __device__ __forceinline__ float2 unpack2(uint32_t u) {
    return __half22float2(*reinterpret_cast<__half2*>(&u));
  }
};
template <> struct Cvt<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
  static __device__ __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  static __device__ __forceinline__ float2 unpack2(uint32_t u) {
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
  }
};
template <> struct Cvt<float> {
  static __device__ __forceinline__ float to_f(float v) { return v; }
  static __device__ __forceinline__ float from_f(float v) { return v; }
};

// ----------------------------------------------------------------------------------------------
// math shared by the fused epilogues (reference: bert_modeling.py:104-116 -- note the constant)
// ----------------------------------------------------------------------------------------------
constexpr float kGeluDiv = 1.41421f;  // the reference divides by 1.41421, not sqrt(2)
constexpr float kGeluInv = 1.0f / 1.41421f;
// Branch-free, MUFU-free erf: odd minimax polynomial (degree 17) on |x| <= 3, clamped to +-1 beyond
// (erf(3) = 0.99998). Max abs error 2.2e-5 -- an order of magnitude below the resolution of the 16-bit
// outputs it feeds. libdevice's erff has a data-dependent branch and exp/rcp MUFU work, which made the
// fused GELU epilogue (64 elements per thread per chunk, two warps per scheduler) issue-bound.
__device__ __forceinline__ float erf_poly(float x) {
  const float xc = fminf(fmaxf(x, -3.0f), 3.0f);
  const float t = xc * xc;
  float p = 4.0746968943e-08f;
  p = fmaf(p, t, -1.9449973089e-06f);
  p = fmaf(p, t, 4.1063089266e-05f);
  p = fmaf(p, t, -5.1105671781e-04f);
  p = fmaf(p, t, 4.2355141407e-03f);
  p = fmaf(p, t, -2.5103073778e-02f);
  p = fmaf(p, t, 1.1107960747e-01f);
  p = fmaf(p, t, -3.7531502903e-01f);
  p = fmaf(p, t, 1.1282684529e+00f);
  return fminf(fmaxf(p * xc, -1.0f), 1.0f);
}
__device__ __forceinline__ float gelu_ref(float x) {
  return x * 0.5f * (1.0f + erf_poly(x * kGeluInv));
}
__device__ __forceinline__ float gelu_grad_ref(float x) {
  const float u = x * kGeluInv;
  // d/dx [0.5 x (1 + erf(x/c))] = 0.5 (1 + erf(u)) + x * exp(-u^2) / (c * sqrt(pi))
  return 0.5f * (1.0f + erf_poly(u)) + x * __expf(-u * u) * (0.5641895835477563f * kGeluInv);
}

// gelu(x) and gelu'(x) together (they share the erf): forward can store the derivative for backward.
// This is the epilogue of the largest GEMM of a layer (64 elements per thread per chunk), so the arithmetic is
// folded as far as it goes -- 19 instructions per element instead of 30: the polynomial takes x itself
// (coefficients of 0.5 * erf(x / 1.41421) = x * Q(x^2), so the scaling of the argument, the "1 +" and the "0.5 *"
// disappear into Q and one final add), and exp(-(x / 1.41421)^2) is one multiply + ex2.approx.ftz (no
// denormal-range fix-up: below 2^-126 the term is 0 in the 16-bit output anyway).
namespace gelu_q {
constexpr double S = 1.0 / 1.41421, S2 = S * S;
constexpr double H1 = 0.5 * S, H3 = H1 * S2, H5 = H3 * S2, H7 = H5 * S2, H9 = H7 * S2, H11 = H9 * S2, H13 = H11 * S2,
                 H15 = H13 * S2, H17 = H15 * S2;
constexpr float Q0 = static_cast<float>(H1 * 1.1282684529e+00), Q1 = static_cast<float>(H3 * -3.7531502903e-01),
                Q2 = static_cast<float>(H5 * 1.1107960747e-01), Q3 = static_cast<float>(H7 * -2.5103073778e-02),
                Q4 = static_cast<float>(H9 * 4.2355141407e-03), Q5 = static_cast<float>(H11 * -5.1105671781e-04),
                Q6 = static_cast<float>(H13 * 4.1063089266e-05), Q7 = static_cast<float>(H15 * -1.9449973089e-06),
                Q8 = static_cast<float>(H17 * 4.0746968943e-08);
constexpr float XMAX = static_cast<float>(3.0 / S);                       // |x / 1.41421| <= 3
constexpr float EXP2_SCALE = static_cast<float>(-S2 * 1.4426950408889634);   // exp(-(x S)^2) = 2^(x^2 * EXP2_SCALE)
constexpr float PDF_SCALE = static_cast<float>(0.5641895835477563 * S);    // d/dx of 0.5 erf(x S) = PDF_SCALE * exp(..)
}  // namespace gelu_q
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void gelu_and_grad(const float x_in, float& y, float& dy) {
  using namespace gelu_q;
  const float x = x_in;     // y may alias the argument
  const float xc = fminf(fmaxf(x, -XMAX), XMAX);
  const float t = xc * xc;
  float q = Q8;
  q = fmaf(q, t, Q7);
  q = fmaf(q, t, Q6);
  q = fmaf(q, t, Q5);
  q = fmaf(q, t, Q4);
  q = fmaf(q, t, Q3);
  q = fmaf(q, t, Q2);
  q = fmaf(q, t, Q1);
  q = fmaf(q, t, Q0);
  const float cdf = fmaf(q, xc, 0.5f);    // |q * xc| <= 0.4999987 on the clamped domain (checked over a 1e-5 grid): no second clamp
  y = x * cdf;
  dy = fmaf(x * ex2_ftz(x * x * EXP2_SCALE), PDF_SCALE, cdf);
}

// ----------------------------------------------------------------------------------------------
// Philox4x32-10 counter RNG: dropout masks are recomputed in backward from (seed, stream, index)
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                            uint32_t k1) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
// 4 uniform floats in [0,1) for element group `idx4` (= element index / 4) of logical stream `stream`.
__device__ __forceinline__ float4 philox_uniform4(uint64_t seed, uint32_t stream, uint64_t idx4) {
  uint4 r = philox4x32(static_cast<uint32_t>(idx4), static_cast<uint32_t>(idx4 >> 32), stream, 0u,
                       static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
  const float s = 2.3283064365386963e-10f;  // 2^-32
  return make_float4(r.x * s, r.y * s, r.z * s, r.w * s);
}

// Dropout keep/scale factors for the 8 consecutive elements of group `idx8` (= element index / 8) of logical
// stream `stream`: ONE Philox call gives 128 bits = 8 x 16-bit uniforms; an element is dropped when its field
// is below thr16 = round(p * 65536) (p is honoured to 1.5e-5). Forward and backward kernels call this with
// the same arguments, so no mask is ever stored.
__device__ __forceinline__ uint64_t philox_seed(const long long* seed_ptr) {
  return static_cast<uint64_t>(seed_ptr[0]) * 0x9E3779B97F4A7C15ull + static_cast<uint64_t>(seed_ptr[1]);
}
__device__ __forceinline__ uint32_t dropout_thr16(float p) { return static_cast<uint32_t>(p * 65536.0f + 0.5f); }
__device__ __forceinline__ void philox_keep8(uint64_t seed, uint32_t stream, uint64_t idx8, uint32_t thr16, float scale,
                                             float (&mk)[8]) {
  const uint4 r = philox4x32(static_cast<uint32_t>(idx8), static_cast<uint32_t>(idx8 >> 32), stream, 0u,
                             static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
  mk[0] = (r.x & 0xFFFFu) >= thr16 ? scale : 0.f; mk[1] = (r.x >> 16) >= thr16 ? scale : 0.f;
  mk[2] = (r.y & 0xFFFFu) >= thr16 ? scale : 0.f; mk[3] = (r.y >> 16) >= thr16 ? scale : 0.f;
  mk[4] = (r.z & 0xFFFFu) >= thr16 ? scale : 0.f; mk[5] = (r.z >> 16) >= thr16 ? scale : 0.f;
  mk[6] = (r.w & 0xFFFFu) >= thr16 ? scale : 0.f; mk[7] = (r.w >> 16) >= thr16 ? scale : 0.f;
}

// ----------------------------------------------------------------------------------------------
// reductions
// ----------------------------------------------------------------------------------------------
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

// Column sums over the 32 lanes of a warp when every lane holds one ROW of N values (N = 32 * 2^k):
// a halving butterfly -- at each step a lane keeps one half of its values and trades the other half with
// its partner -- so N-32 + ... shuffles instead of 5*N. Afterwards v[0 .. N/32) of lane L hold the sums of
// columns L*(N/32) .. L*(N/32) + N/32 - 1. Destroys v.
template <int N>
__device__ __forceinline__ void warp_colsum(float (&v)[N], int lane) {
  static_assert(N >= 32 && (N & (N - 1)) == 0, "N must be a power of two >= 32");
  int half = N / 2;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
      if (i < half) {
        const float send = up ? v[i] : v[i + half];
        const float keep = up ? v[i + half] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
      }
    }
    half >>= 1;
  }
}

}  // namespace hs
