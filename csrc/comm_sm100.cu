// hetseq_b200 -- the fused data-parallel gradient path over NVLink peer memory (no NCCL call).
//
// The reference reduces gradients with DDP's bucketed NCCL all-reduce (controller.py:78-86; SURVEY.md
// N4) and then runs multiply_grads, clip_grad_norm and Adam as per-tensor loops on EVERY rank over ALL
// parameters (optim.py:59-70,162-231; K15-K17). Here W ranks of one NVSwitch domain map each other's flat
// gradient / parameter buffers (CUDA IPC) and four kernels do the whole job on each rank:
//
//   hs_dp_reduce_scatter   per bucket, overlapped with backward: announce "my bucket is final" in every
//                          peer's signal pad, wait for all peers, then for each OWNED 2048-element chunk
//                          load the W copies straight from peer memory (ld.relaxed.sys over NVLink), add
//                          them, keep the fp32 sum locally and fold its square into the norm partial.
//   hs_dp_norm_exchange    W floats cross the fabric: every rank deposits its partial in every pad, all
//                          ranks form the same global norm, clip coefficient and overflow flag.
//   hs_dp_adam_allgather   fused unscale * clip * Adam on the owned chunks (1/W of the state and of the
//                          HBM traffic), the new 16-bit parameters are stored directly into ALL ranks'
//                          parameter buffers (st over NVLink) -- the all-gather is the update's epilogue.
//   hs_dp_wait_params      returns once every peer's parameter writes for this step have landed.
//
// Ownership is cyclic by chunk (chunk c belongs to rank c % W), so any bucket spreads evenly over all
// GPUs' NVLink ingress. Bytes on the wire per GPU per step: (W-1)/W * N fp32 in, (W-1)/W * N 16-bit out,
// versus 2 * (W-1)/W * N each way for a ring all-reduce. Cross-GPU ordering: release/acquire at .sys
// scope on epoch-numbered flags (never reset), data loads/stores at .sys scope; every spin is bounded and
// traps instead of hanging the GPU if a peer disappears.
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

#include <cstring>

namespace hs {

constexpr int DP_THREADS = 512;
// The reduce-scatter runs WHILE backward's GEMMs own the SMs (320 threads x 168 registers, ~225 KB smem per
// SM). It only overlaps if its CTAs fit into what is left of an SM: no shared memory to speak of and
// <= 11.7 K registers. 256 threads at <= 42 registers (launch bound 6 CTAs/SM) is 10.5 K.
constexpr int RS_THREADS = 256;
constexpr int DP_MAX_WORLD = 16;

// signal pad layout (32-bit words)
__host__ __device__ inline int sig_param_ready(int src) { return src; }
__host__ __device__ inline int sig_norm_flag(int W, int src) { return W + src; }
__host__ __device__ inline int sig_norm_value(int W, int src) { return 2 * W + src; }
__host__ __device__ inline int sig_overflow_value(int W, int src) { return 3 * W + src; }
__host__ __device__ inline int sig_grad_ready(int W, int slot, int src) { return 4 * W + slot * W + src; }

struct PeerTable {
  void* ptr[DP_MAX_WORLD];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
// Streaming variants for the reduce-scatter: 1.3 GB of gradients pass through once. Without a hint they are
// allocated in L2 with normal priority -- on the reading GPU for its own copy and on the serving GPU for peer
// reads -- and evict the activation / weight tiles that backward's GEMMs re-read from L2 (measured on 2 GPUs:
// backward 1.0 ms slower while the reduce kernels ran next to it). evict_first keeps them out of the way.
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ float4 ld_peer_stream_f4(const float4* p, uint64_t pol) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p), "l"(pol)
               : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_local_stream_f4(const float4* p, uint64_t pol) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p), "l"(pol)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_local_stream_f4(float4* p, float4 v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;"
               ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_u2(uint2* p, uint2 v) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// Spin until *p >= want (epochs only grow). Bounded: a vanished peer becomes a trap, not a hung GPU.
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t want, const char* what, int peer) {
  if (static_cast<int32_t>(ld_acquire_sys(p) - want) >= 0) return;
  const long long t0 = clock64();
  while (static_cast<int32_t>(ld_acquire_sys(p) - want) < 0) {
    __nanosleep(64);
    if (clock64() - t0 > 40000000000LL) {
      printf("hetseq_b200: timed out waiting for %s from rank %d (want epoch %u)\n", what, peer, want);
      __trap();
    }
  }
}

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS, 6)
dp_reduce_scatter_kernel(PeerTable grads, PeerTable sigs, int rank, int W, long long chunk_begin,
                         long long chunk_end, int chunk, int slot, uint32_t epoch, const uint32_t* __restrict__ epoch_dev,
                         float* __restrict__ sumsq_accum) {
  __shared__ float red[RS_THREADS / 32];
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  // 1. announce + wait (block 0 announces; every block waits on its own before touching peer data)
  if (blockIdx.x == 0 && threadIdx.x < W) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(sigs.ptr[threadIdx.x]) + sig_grad_ready(W, slot, rank), epoch);
  }
  // our own gradients are final by stream order; only the peers' announcements need waiting for
  if (threadIdx.x < W && threadIdx.x != rank) {
    wait_flag(reinterpret_cast<const uint32_t*>(sigs.ptr[rank]) + sig_grad_ready(W, slot, threadIdx.x), epoch,
              "gradient bucket", threadIdx.x);
  }
  __syncthreads();

  // 2. owned chunks of this bucket: c in [chunk_begin, chunk_end), c % W == rank
  long long first = chunk_begin + ((rank - chunk_begin % W) + W) % W;
  const int vec_per_chunk = chunk / 4;
  float acc = 0.f;
  float* mine = reinterpret_cast<float*>(grads.ptr[rank]);
  const uint64_t pol = l2_evict_first_policy();
  for (long long c = first + static_cast<long long>(blockIdx.x) * W; c < chunk_end; c += static_cast<long long>(gridDim.x) * W) {
    const long long base = c * chunk;
    // two vectors per thread and all peers' loads issued before the first add: (W-1)*2 NVLink loads in flight
    for (int i = threadIdx.x; i < vec_per_chunk; i += 2 * blockDim.x) {
      const int i2 = i + blockDim.x;
      const bool has2 = i2 < vec_per_chunk;
      float4 s0 = ld_local_stream_f4(reinterpret_cast<const float4*>(mine + base) + i, pol);
      float4 s1 = has2 ? ld_local_stream_f4(reinterpret_cast<const float4*>(mine + base) + i2, pol)
                       : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
      for (int p = 1; p < W; ++p) {
        const int peer = (rank + p) % W;   // stagger peers so that ranks do not all hit the same GPU at once
        const float4* src = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(grads.ptr[peer]) + base);
        const float4 v0 = ld_peer_stream_f4(src + i, pol);
        float4 v1 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (has2) v1 = ld_peer_stream_f4(src + i2, pol);
        s0.x += v0.x; s0.y += v0.y; s0.z += v0.z; s0.w += v0.w;
        s1.x += v1.x; s1.y += v1.y; s1.z += v1.z; s1.w += v1.w;
      }
      st_local_stream_f4(reinterpret_cast<float4*>(mine + base) + i, s0, pol);
      acc += s0.x * s0.x + s0.y * s0.y + s0.z * s0.z + s0.w * s0.w;
      if (has2) {
        st_local_stream_f4(reinterpret_cast<float4*>(mine + base) + i2, s1, pol);
        acc += s1.x * s1.x + s1.y * s1.y + s1.z * s1.z + s1.w * s1.w;
      }
    }
  }
  // 3. norm partial
  acc = warp_sum(acc);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) red[warp] = acc;
  __syncthreads();
  if (warp == 0) {
    acc = lane < (blockDim.x >> 5) ? red[lane] : 0.f;
    acc = warp_sum(acc);
    if (lane == 0) atomicAdd(sumsq_accum, acc);
  }
}

// out = [grad_norm, multiplier (incl. clip), skip flag (as float), global sum of squares]
__global__ void __launch_bounds__(32)
dp_norm_exchange_kernel(PeerTable sigs, int rank, int W, uint32_t epoch, const uint32_t* __restrict__ epoch_dev,
                        float* __restrict__ sumsq_accum, float mult_host, const float* __restrict__ mult_dev,
                        float max_norm, float* __restrict__ out, int* __restrict__ skip_out) {
  const int t = threadIdx.x;
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  const float mine = *sumsq_accum;
  if (t < W) {
    uint32_t* pad = reinterpret_cast<uint32_t*>(sigs.ptr[t]);
    pad[sig_norm_value(W, rank)] = __float_as_uint(mine);
    __threadfence_system();
    st_release_sys(pad + sig_norm_flag(W, rank), epoch);
  }
  float v = 0.f;
  if (t < W) {
    const uint32_t* pad = reinterpret_cast<const uint32_t*>(sigs.ptr[rank]);
    wait_flag(pad + sig_norm_flag(W, t), epoch, "norm partial", t);
    v = __uint_as_float(*reinterpret_cast<const volatile uint32_t*>(pad + sig_norm_value(W, t)));
  }
  // same summation order on every rank -> bit-identical norm everywhere
  float total = 0.f;
  for (int p = 0; p < W; ++p) total += __shfl_sync(0xffffffffu, v, p);
  if (t == 0) {
    const float mult = mult_host * (mult_dev != nullptr ? *mult_dev : 1.0f);
    const float norm = sqrtf(total) * fabsf(mult);
    const bool bad = !(norm == norm) || isinf(norm);
    float m = mult;
    if (!bad && max_norm > 0.f) m *= fminf(max_norm / (norm + 1e-6f), 1.0f);
    out[0] = norm;
    out[1] = m;
    out[2] = bad ? 1.f : 0.f;
    out[3] = total;
    *skip_out = bad ? 1 : 0;
    *sumsq_accum = 0.f;   // ready for the next step's accumulation
  }
}

struct DpAdamHyper {
  float lr, beta1, beta2, eps, weight_decay, step_factor;
};

template <typename OutT>
__global__ void __launch_bounds__(DP_THREADS)
dp_adam_allgather_kernel(const float* __restrict__ grad, float* __restrict__ master, float* __restrict__ m,
                         float* __restrict__ v, PeerTable params, PeerTable sigs, int rank, int W,
                         long long nblocks, int chunk, DpAdamHyper h, const float* __restrict__ scalars,
                         const float* __restrict__ hyper_dev, uint32_t epoch, const uint32_t* __restrict__ epoch_dev,
                         unsigned int* __restrict__ done_counter) {
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  const bool skip = scalars[2] != 0.f;
  if (!skip) {
    const float mult = scalars[1];
    const float lr = hyper_dev != nullptr ? hyper_dev[0] : h.lr;
    const float step_size = lr * (hyper_dev != nullptr ? hyper_dev[1] : h.step_factor);
    const float decay = 1.0f - h.weight_decay * lr;
    const float b1 = h.beta1, b2 = h.beta2, ob1 = 1.0f - h.beta1, ob2 = 1.0f - h.beta2;
    const int vec_per_chunk = chunk / 4;
    const long long stride = static_cast<long long>(W) * chunk;
    for (long long c = blockIdx.x; c < nblocks; c += gridDim.x) {
      const long long goff = c * stride + static_cast<long long>(rank) * chunk;   // position in the flat buffers
      const float4* g4 = reinterpret_cast<const float4*>(grad + goff);
      float4* p4 = reinterpret_cast<float4*>(master + c * static_cast<long long>(chunk));
      float4* m4 = reinterpret_cast<float4*>(m + c * static_cast<long long>(chunk));
      float4* v4 = reinterpret_cast<float4*>(v + c * static_cast<long long>(chunk));
      for (int i = threadIdx.x; i < vec_per_chunk; i += blockDim.x) {
        const float4 g = g4[i];
        float4 mm = m4[i], vv = v4[i], pp = p4[i];
        float gs[4] = {g.x * mult, g.y * mult, g.z * mult, g.w * mult};
        float ms[4] = {mm.x, mm.y, mm.z, mm.w}, vs[4] = {vv.x, vv.y, vv.z, vv.w}, ps[4] = {pp.x, pp.y, pp.z, pp.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          ms[k] = b1 * ms[k] + ob1 * gs[k];
          vs[k] = b2 * vs[k] + ob2 * gs[k] * gs[k];
          ps[k] = ps[k] * decay - step_size * (ms[k] / (sqrtf(vs[k]) + h.eps));
        }
        m4[i] = make_float4(ms[0], ms[1], ms[2], ms[3]);
        v4[i] = make_float4(vs[0], vs[1], vs[2], vs[3]);
        p4[i] = make_float4(ps[0], ps[1], ps[2], ps[3]);
        uint2 o;
        o.x = Cvt<OutT>::pack2(ps[0], ps[1]);
        o.y = Cvt<OutT>::pack2(ps[2], ps[3]);
        // the all-gather: the update's epilogue stores the new parameters into every rank's buffer
#pragma unroll 4
        for (int p = 0; p < W; ++p) {
          const int peer = (rank + p) % W;
          st_relaxed_sys_u2(reinterpret_cast<uint2*>(reinterpret_cast<OutT*>(params.ptr[peer]) + goff) + i, o);
        }
      }
    }
  }
  // last block to finish publishes "my parameters have landed" to every rank
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done_counter, 1u);
    last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *done_counter = 0;
    if (threadIdx.x < W) {
      __threadfence_system();
      st_release_sys(reinterpret_cast<uint32_t*>(sigs.ptr[threadIdx.x]) + sig_param_ready(rank), epoch);
    }
  }
}

__global__ void __launch_bounds__(32)
dp_wait_params_kernel(PeerTable sigs, int rank, int W, uint32_t epoch, const uint32_t* __restrict__ epoch_dev) {
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  if (threadIdx.x < W)
    wait_flag(reinterpret_cast<const uint32_t*>(sigs.ptr[rank]) + sig_param_ready(threadIdx.x), epoch,
              "updated parameters", threadIdx.x);
}

static int fill_table(PeerTable* t, void* const* ptrs, int W) {
  if (W < 1 || W > DP_MAX_WORLD) { set_last_error("fused data-parallel path supports 1..%d ranks (got %d)", DP_MAX_WORLD, W); return 1; }
  for (int i = 0; i < DP_MAX_WORLD; ++i) t->ptr[i] = i < W ? ptrs[i] : nullptr;
  return 0;
}

}  // namespace hs

using namespace hs;

// ---- symmetric memory plumbing (CUDA IPC) ---------------------------------------------------------
extern "C" int hs_ipc_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
  void* p = nullptr;
  HS_CUDA_OK(cudaMalloc(&p, bytes));
  HS_CUDA_OK(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  HS_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  *ptr = p;
  return 0;
}
extern "C" int hs_ipc_open(const unsigned char* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  HS_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int hs_ipc_close(void* ptr) { HS_CUDA_OK(cudaIpcCloseMemHandle(ptr)); return 0; }
extern "C" int hs_ipc_free(void* ptr) { HS_CUDA_OK(cudaFree(ptr)); return 0; }
extern "C" int hs_dp_signal_words(int W, int max_buckets) { return 4 * W + max_buckets * W; }

// ---- kernels ----------------------------------------------------------------------------------------
extern "C" int hs_dp_reduce_scatter(void* const* grad_ptrs, void* const* sig_ptrs, int rank, int W,
                                    long long chunk_begin, long long chunk_end, int chunk, int slot, unsigned epoch,
                                    const unsigned* epoch_dev, float* sumsq_accum, int nblocks, hs_stream_t stream) {
  PeerTable g, s;
  if (fill_table(&g, grad_ptrs, W) || fill_table(&s, sig_ptrs, W)) return 1;
  if (chunk % 4 != 0 || chunk_end <= chunk_begin || rank < 0 || rank >= W) {
    set_last_error("hs_dp_reduce_scatter: bad arguments rank=%d W=%d chunks=[%lld,%lld) chunk=%d", rank, W, chunk_begin, chunk_end, chunk);
    return 1;
  }
  // nblocks <= 0: one CTA per owned chunk (short-lived CTAs). What decides how well this kernel overlaps with
  // backward turned out to be stream PRIORITY, not CTA lifetime: launched at a higher priority than the compute
  // stream its pending CTAs took every SM slot freed at a GEMM kernel boundary and the next GEMM's CTAs (which
  // need almost a whole SM each) waited; at a lower priority it only uses what backward leaves free.
  const long long owned = (chunk_end - chunk_begin + W - 1) / W;
  long long want = nblocks > 0 ? nblocks : owned;
  if (want > owned) want = owned;
  if (want > 1 << 20) want = 1 << 20;
  const int grid = static_cast<int>(want < 1 ? 1 : want);
  dp_reduce_scatter_kernel<<<grid, RS_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(g, s, rank, W, chunk_begin, chunk_end, chunk, slot, epoch, epoch_dev, sumsq_accum);
  return post_launch("dp_reduce_scatter_kernel");
}

extern "C" int hs_dp_norm_exchange(void* const* sig_ptrs, int rank, int W, unsigned epoch, const unsigned* epoch_dev,
                                   float* sumsq_accum,
                                   float mult_host, const float* mult_dev, float max_norm, float* out4, int* skip_out,
                                   hs_stream_t stream) {
  PeerTable s;
  if (fill_table(&s, sig_ptrs, W)) return 1;
  dp_norm_exchange_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(s, rank, W, epoch, epoch_dev, sumsq_accum, mult_host, mult_dev, max_norm, out4, skip_out);
  return post_launch("dp_norm_exchange_kernel");
}

extern "C" int hs_dp_adam_allgather(const float* grad, float* master, float* exp_avg, float* exp_avg_sq,
                                    void* const* param_ptrs, void* const* sig_ptrs, int rank, int W, long long nblocks,
                                    int chunk, int out_dtype, float lr, float beta1, float beta2, float eps,
                                    float weight_decay, float step_factor, const float* scalars, const float* hyper_dev,
                                    unsigned epoch, const unsigned* epoch_dev, unsigned int* done_counter, int grid_blocks,
                                    hs_stream_t stream) {
  PeerTable p, s;
  if (fill_table(&p, param_ptrs, W) || fill_table(&s, sig_ptrs, W)) return 1;
  if (chunk % 4 != 0 || nblocks <= 0) { set_last_error("hs_dp_adam_allgather: bad layout"); return 1; }
  DpAdamHyper h{lr, beta1, beta2, eps, weight_decay, step_factor};
  int grid = grid_blocks > 0 ? grid_blocks : num_sms() * 2;
  if (grid > nblocks) grid = static_cast<int>(nblocks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_dtype == HS_F16)
    dp_adam_allgather_kernel<__half><<<grid, DP_THREADS, 0, st>>>(grad, master, exp_avg, exp_avg_sq, p, s, rank, W, nblocks, chunk, h, scalars, hyper_dev, epoch, epoch_dev, done_counter);
  else if (out_dtype == HS_BF16)
    dp_adam_allgather_kernel<__nv_bfloat16><<<grid, DP_THREADS, 0, st>>>(grad, master, exp_avg, exp_avg_sq, p, s, rank, W, nblocks, chunk, h, scalars, hyper_dev, epoch, epoch_dev, done_counter);
  else { set_last_error("hs_dp_adam_allgather: parameters must be f16 or bf16"); return 1; }
  return post_launch("dp_adam_allgather_kernel");
}

extern "C" int hs_dp_wait_params(void* const* sig_ptrs, int rank, int W, unsigned epoch, const unsigned* epoch_dev,
                                 hs_stream_t stream) {
  PeerTable s;
  if (fill_table(&s, sig_ptrs, W)) return 1;
  dp_wait_params_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(s, rank, W, epoch, epoch_dev);
  return post_launch("dp_wait_params_kernel");
}
