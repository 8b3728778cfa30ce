// hetseq_b200 -- flat-buffer gradient norm and fused optimizer updates.
//
// Replaces the reference's per-tensor Python loops: multiply_grads (optim.py:59-63), clip_grad_norm
// (optim.py:65-70) and Adam.step / Adadelta.step (optim.py:162-231, 263-304) -- ~10 elementwise
// launches per tensor, 2-4 k launches per step for BERT -- with
//   hs_sumsq        : sum of squares of the owned gradient elements (two deterministic stages)
//   hs_adam_flat    : grad * mult -> m, v -> decoupled decay -> update -> fp32 master + 16-bit param
//   hs_adadelta_flat: same structure for Adadelta
// "Owned" elements are described by (nchunks, chunk, stride): chunk c covers elements
// [c*stride, c*stride + chunk) of the flat gradient / parameter buffers, while optimizer state and
// master weights are compact (c*chunk + i). Unsharded: nchunks = total/chunk, stride = chunk.
// Sharded over W ranks: stride = W*chunk and the base pointers are offset by rank*chunk
// (parallel/flat_buffer.py). All hot loops are 128-bit vectorised and purely HBM-bound.
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

#include <mutex>

namespace hs {

constexpr int OPT_THREADS = 256;

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  if (warp == 0) {
    v = (lane < (blockDim.x >> 5)) ? red[lane] : 0.0f;
    v = warp_sum(v);
  }
  return v;  // valid in thread 0
}

template <typename T>
__global__ void __launch_bounds__(OPT_THREADS)
sumsq_partial_kernel(const T* __restrict__ g, long long nchunks, int chunk, long long stride,
                     float* __restrict__ partials) {
  __shared__ float red[32];
  float acc = 0.0f;
  constexpr int VEC = 16 / sizeof(T);
  const int vec_per_chunk = chunk / VEC;
  for (long long c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const uint4* src = reinterpret_cast<const uint4*>(g + c * stride);
    for (int i = threadIdx.x; i < vec_per_chunk; i += blockDim.x) {
      const uint4 u = src[i];
      if constexpr (sizeof(T) == 4) {
        const float a = __uint_as_float(u.x), b = __uint_as_float(u.y), cc = __uint_as_float(u.z),
                    d = __uint_as_float(u.w);
        acc += a * a + b * b + cc * cc + d * d;
      } else {
        const float2 f0 = Cvt<T>::unpack2(u.x), f1 = Cvt<T>::unpack2(u.y), f2 = Cvt<T>::unpack2(u.z),
                     f3 = Cvt<T>::unpack2(u.w);
        acc += f0.x * f0.x + f0.y * f0.y + f1.x * f1.x + f1.y * f1.y + f2.x * f2.x + f2.y * f2.y +
               f3.x * f3.x + f3.y * f3.y;
      }
    }
  }
  const float s = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = s;
}

__global__ void __launch_bounds__(OPT_THREADS)
sumsq_final_kernel(const float* __restrict__ partials, int n, float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  float acc = 0.0f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += partials[i];
  const float s = block_sum(acc, red);
  if (threadIdx.x == 0) out[0] = accumulate ? out[0] + s : s;
}

struct AdamHyper {
  float lr, beta1, beta2, eps, weight_decay, step_factor;  // step_factor = sqrt(1-b2^t)/(1-b1^t)
};

// hyper_dev (optional): [lr, step_factor] read from device memory (keeps captured graphs valid)
// mult_dev  (optional): gradient multiplier computed on the device (clip coefficient etc.)
// skip_dev  (optional): non-zero -> leave everything untouched (fp16 overflow step)
template <typename OutT>
__global__ void __launch_bounds__(OPT_THREADS)
adam_flat_kernel(const float* __restrict__ grad, float* __restrict__ master, float* __restrict__ m,
                 float* __restrict__ v, OutT* __restrict__ param_out, long long nchunks, int chunk,
                 long long stride, AdamHyper h, float mult_host, const float* __restrict__ mult_dev,
                 const float* __restrict__ hyper_dev, const int* __restrict__ skip_dev) {
  if (skip_dev != nullptr && *skip_dev != 0) return;
  const float mult = mult_host * (mult_dev != nullptr ? *mult_dev : 1.0f);
  const float lr = hyper_dev != nullptr ? hyper_dev[0] : h.lr;
  const float step_size = lr * (hyper_dev != nullptr ? hyper_dev[1] : h.step_factor);
  const float decay = 1.0f - h.weight_decay * lr;
  const float b1 = h.beta1, b2 = h.beta2, ob1 = 1.0f - h.beta1, ob2 = 1.0f - h.beta2;
  const int vec_per_chunk = chunk / 4;
  const bool alias = reinterpret_cast<const void*>(param_out) == reinterpret_cast<const void*>(master);
  for (long long c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const float4* g4 = reinterpret_cast<const float4*>(grad + c * stride);
    float4* p4 = reinterpret_cast<float4*>(master + c * (long long)chunk);
    float4* m4 = reinterpret_cast<float4*>(m + c * (long long)chunk);
    float4* v4 = reinterpret_cast<float4*>(v + c * (long long)chunk);
    OutT* out = param_out + c * stride;
    for (int i = threadIdx.x; i < vec_per_chunk; i += blockDim.x) {
      const float4 g = g4[i];
      float4 mm = m4[i], vv = v4[i], pp = p4[i];
      float gs[4] = {g.x * mult, g.y * mult, g.z * mult, g.w * mult};
      float ms[4] = {mm.x, mm.y, mm.z, mm.w}, vs[4] = {vv.x, vv.y, vv.z, vv.w}, ps[4] = {pp.x, pp.y, pp.z, pp.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        ms[k] = b1 * ms[k] + ob1 * gs[k];
        vs[k] = b2 * vs[k] + ob2 * gs[k] * gs[k];
        const float denom = sqrtf(vs[k]) + h.eps;
        ps[k] = ps[k] * decay - step_size * (ms[k] / denom);   // decay first, then the Adam update
      }
      m4[i] = make_float4(ms[0], ms[1], ms[2], ms[3]);
      v4[i] = make_float4(vs[0], vs[1], vs[2], vs[3]);
      p4[i] = make_float4(ps[0], ps[1], ps[2], ps[3]);
      if constexpr (sizeof(OutT) == 2) {
        uint2 o;
        o.x = Cvt<OutT>::pack2(ps[0], ps[1]);
        o.y = Cvt<OutT>::pack2(ps[2], ps[3]);
        reinterpret_cast<uint2*>(out)[i] = o;
      } else {
        if (!alias)   // fp32 model: master aliases the parameters
          reinterpret_cast<float4*>(out)[i] = make_float4(ps[0], ps[1], ps[2], ps[3]);
      }
    }
  }
}

template <typename OutT>
__global__ void __launch_bounds__(OPT_THREADS)
adadelta_flat_kernel(const float* __restrict__ grad, float* __restrict__ master, float* __restrict__ sq,
                     float* __restrict__ acc, OutT* __restrict__ param_out, long long nchunks, int chunk,
                     long long stride, float lr, float rho, float eps, float weight_decay, float mult_host,
                     const float* __restrict__ mult_dev, const float* __restrict__ hyper_dev,
                     const int* __restrict__ skip_dev) {
  if (skip_dev != nullptr && *skip_dev != 0) return;
  const float mult = mult_host * (mult_dev != nullptr ? *mult_dev : 1.0f);
  if (hyper_dev != nullptr) lr = hyper_dev[0];
  const int vec_per_chunk = chunk / 4;
  const bool alias = reinterpret_cast<const void*>(param_out) == reinterpret_cast<const void*>(master);
  for (long long c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const float4* g4 = reinterpret_cast<const float4*>(grad + c * stride);
    float4* p4 = reinterpret_cast<float4*>(master + c * (long long)chunk);
    float4* s4 = reinterpret_cast<float4*>(sq + c * (long long)chunk);
    float4* a4 = reinterpret_cast<float4*>(acc + c * (long long)chunk);
    OutT* out = param_out + c * stride;
    for (int i = threadIdx.x; i < vec_per_chunk; i += blockDim.x) {
      const float4 g = g4[i];
      float4 ss = s4[i], aa = a4[i], pp = p4[i];
      float gs[4] = {g.x * mult, g.y * mult, g.z * mult, g.w * mult};
      float sv[4] = {ss.x, ss.y, ss.z, ss.w}, av[4] = {aa.x, aa.y, aa.z, aa.w}, ps[4] = {pp.x, pp.y, pp.z, pp.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float gk = gs[k] + weight_decay * ps[k];           // coupled L2 decay (optim.py:295-296)
        sv[k] = rho * sv[k] + (1.0f - rho) * gk * gk;
        const float delta = sqrtf(av[k] + eps) / sqrtf(sv[k] + eps) * gk;
        ps[k] -= lr * delta;
        av[k] = rho * av[k] + (1.0f - rho) * delta * delta;
      }
      s4[i] = make_float4(sv[0], sv[1], sv[2], sv[3]);
      a4[i] = make_float4(av[0], av[1], av[2], av[3]);
      p4[i] = make_float4(ps[0], ps[1], ps[2], ps[3]);
      if constexpr (sizeof(OutT) == 2) {
        uint2 o;
        o.x = Cvt<OutT>::pack2(ps[0], ps[1]);
        o.y = Cvt<OutT>::pack2(ps[2], ps[3]);
        reinterpret_cast<uint2*>(out)[i] = o;
      } else {
        if (!alias)
          reinterpret_cast<float4*>(out)[i] = make_float4(ps[0], ps[1], ps[2], ps[3]);
      }
    }
  }
}

static int opt_grid(long long nchunks) {
  const long long cap = (long long)num_sms() * 8;
  return (int)(nchunks < cap ? (nchunks < 1 ? 1 : nchunks) : cap);
}

static int check_layout(const char* who, long long nchunks, int chunk, long long stride) {
  if (nchunks <= 0 || chunk <= 0 || chunk % 8 != 0 || stride < chunk || stride % 8 != 0) {
    set_last_error("%s: bad layout nchunks=%lld chunk=%d stride=%lld (chunk and stride must be multiples of 8)",
                   who, nchunks, chunk, stride);
    return 1;
  }
  return 0;
}

}  // namespace hs

using namespace hs;

extern "C" int hs_sumsq_scratch_elems(void) { return num_sms() * 8; }

extern "C" int hs_sumsq(const void* grad, int dtype, long long nchunks, int chunk, long long stride,
                        float* scratch, float* out, int accumulate, hs_stream_t stream) {
  if (check_layout("hs_sumsq", nchunks, chunk, stride)) return 1;
  if (scratch == nullptr || out == nullptr) { set_last_error("hs_sumsq: scratch and out are required"); return 1; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = opt_grid(nchunks);
  if (dtype == HS_F32) sumsq_partial_kernel<float><<<grid, OPT_THREADS, 0, s>>>((const float*)grad, nchunks, chunk, stride, scratch);
  else if (dtype == HS_F16) sumsq_partial_kernel<__half><<<grid, OPT_THREADS, 0, s>>>((const __half*)grad, nchunks, chunk, stride, scratch);
  else if (dtype == HS_BF16) sumsq_partial_kernel<__nv_bfloat16><<<grid, OPT_THREADS, 0, s>>>((const __nv_bfloat16*)grad, nchunks, chunk, stride, scratch);
  else { set_last_error("hs_sumsq: bad dtype %d", dtype); return 1; }
  if (post_launch("sumsq_partial_kernel")) return 1;
  sumsq_final_kernel<<<1, OPT_THREADS, 0, s>>>(scratch, grid, out, accumulate);
  return post_launch("sumsq_final_kernel");
}

// ---- allocation-free variant of hs_sumsq ----------------------------------------------------------
// Some hosts of this library cannot hand it a scratch buffer: a process that has given the whole GPU
// to its own memory pool up front (any later cudaMalloc may fail, and cudaMalloc/cudaFree synchronise
// the device, which a latency-critical step loop cannot afford), or a plain C/C++ program that links
// the C ABI and has no caching allocator to borrow from. For them the block partials live in a small
// array that is part of the library image itself. Every GPU that loads the library holds these 16 KiB
// until the process exits.
constexpr int HS_SUMSQ_NOALLOC_MAX_PARTIALS = 4096;
static __device__ __align__(256) float g_sumsq_static_scratch[HS_SUMSQ_NOALLOC_MAX_PARTIALS];

namespace hs {
// In host code the symbol above is only a handle; the device address has to be looked up, once per
// device and with that device current.
static int sumsq_static_scratch(float** out) {
  static float* ptr_by_device[64] = {nullptr};
  static std::mutex mu;
  int dev = 0;
  HS_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { set_last_error("hs_sumsq_noalloc: device index %d is out of range", dev); return 1; }
  std::lock_guard<std::mutex> lock(mu);
  if (ptr_by_device[dev] == nullptr) {
    float* p = nullptr;
    HS_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_sumsq_static_scratch));
    ptr_by_device[dev] = p;
  }
  *out = ptr_by_device[dev];
  return 0;
}
}  // namespace hs

extern "C" int hs_sumsq_noalloc_max_partials(void) { return HS_SUMSQ_NOALLOC_MAX_PARTIALS; }

extern "C" int hs_sumsq_noalloc(const void* grad, int dtype, long long nchunks, int chunk, long long stride,
                                float* out, int accumulate, hs_stream_t stream) {
  if (check_layout("hs_sumsq_noalloc", nchunks, chunk, stride)) return 1;
  if (out == nullptr) { set_last_error("hs_sumsq_noalloc: out is required"); return 1; }
  const int grid = opt_grid(nchunks);
  if (grid > HS_SUMSQ_NOALLOC_MAX_PARTIALS) {
    set_last_error("hs_sumsq_noalloc: this launch needs %d block partials but the built-in scratch holds %d; "
                   "use hs_sumsq with caller-provided scratch", grid, HS_SUMSQ_NOALLOC_MAX_PARTIALS);
    return 1;
  }
  float* scratch = nullptr;
  if (sumsq_static_scratch(&scratch)) return 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // Stage 1 overwrites partials [0, grid) and stage 2 reads exactly those, so nothing left behind by an
  // earlier call is ever consumed and nothing needs clearing first.
  if (dtype == HS_F32) sumsq_partial_kernel<float><<<grid, OPT_THREADS, 0, s>>>((const float*)grad, nchunks, chunk, stride, scratch);
  else if (dtype == HS_F16) sumsq_partial_kernel<__half><<<grid, OPT_THREADS, 0, s>>>((const __half*)grad, nchunks, chunk, stride, scratch);
  else if (dtype == HS_BF16) sumsq_partial_kernel<__nv_bfloat16><<<grid, OPT_THREADS, 0, s>>>((const __nv_bfloat16*)grad, nchunks, chunk, stride, scratch);
  else { set_last_error("hs_sumsq_noalloc: bad dtype %d", dtype); return 1; }
  if (post_launch("sumsq_partial_kernel")) return 1;
  sumsq_final_kernel<<<1, OPT_THREADS, 0, s>>>(scratch, grid, out, accumulate);
  return post_launch("sumsq_final_kernel");
}

extern "C" int hs_adam_flat(const float* grad, float* master, float* exp_avg, float* exp_avg_sq, void* param_out,
                            int out_dtype, long long nchunks, int chunk, long long stride, float lr, float beta1,
                            float beta2, float eps, float weight_decay, float step_factor, float mult_host,
                            const float* mult_dev, const float* hyper_dev, const int* skip_dev,
                            hs_stream_t stream) {
  if (check_layout("hs_adam_flat", nchunks, chunk, stride)) return 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, step_factor};
  const int grid = opt_grid(nchunks);
  if (out_dtype == HS_F16)
    adam_flat_kernel<__half><<<grid, OPT_THREADS, 0, s>>>(grad, master, exp_avg, exp_avg_sq, (__half*)param_out, nchunks, chunk, stride, h, mult_host, mult_dev, hyper_dev, skip_dev);
  else if (out_dtype == HS_BF16)
    adam_flat_kernel<__nv_bfloat16><<<grid, OPT_THREADS, 0, s>>>(grad, master, exp_avg, exp_avg_sq, (__nv_bfloat16*)param_out, nchunks, chunk, stride, h, mult_host, mult_dev, hyper_dev, skip_dev);
  else if (out_dtype == HS_F32)
    adam_flat_kernel<float><<<grid, OPT_THREADS, 0, s>>>(grad, master, exp_avg, exp_avg_sq, (float*)param_out, nchunks, chunk, stride, h, mult_host, mult_dev, hyper_dev, skip_dev);
  else { set_last_error("hs_adam_flat: bad out dtype %d", out_dtype); return 1; }
  return post_launch("adam_flat_kernel");
}

extern "C" int hs_adadelta_flat(const float* grad, float* master, float* square_avg, float* acc_delta,
                                void* param_out, int out_dtype, long long nchunks, int chunk, long long stride,
                                float lr, float rho, float eps, float weight_decay, float mult_host,
                                const float* mult_dev, const float* hyper_dev, const int* skip_dev,
                                hs_stream_t stream) {
  if (check_layout("hs_adadelta_flat", nchunks, chunk, stride)) return 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = opt_grid(nchunks);
  if (out_dtype == HS_F16)
    adadelta_flat_kernel<__half><<<grid, OPT_THREADS, 0, s>>>(grad, master, square_avg, acc_delta, (__half*)param_out, nchunks, chunk, stride, lr, rho, eps, weight_decay, mult_host, mult_dev, hyper_dev, skip_dev);
  else if (out_dtype == HS_BF16)
    adadelta_flat_kernel<__nv_bfloat16><<<grid, OPT_THREADS, 0, s>>>(grad, master, square_avg, acc_delta, (__nv_bfloat16*)param_out, nchunks, chunk, stride, lr, rho, eps, weight_decay, mult_host, mult_dev, hyper_dev, skip_dev);
  else if (out_dtype == HS_F32)
    adadelta_flat_kernel<float><<<grid, OPT_THREADS, 0, s>>>(grad, master, square_avg, acc_delta, (float*)param_out, nchunks, chunk, stride, lr, rho, eps, weight_decay, mult_host, mult_dev, hyper_dev, skip_dev);
  else { set_last_error("hs_adadelta_flat: bad out dtype %d", out_dtype); return 1; }
  return post_launch("adadelta_flat_kernel");
}
