// hetseq_b200 -- the fused data-parallel gradient path over NVLink / NVSwitch (no NCCL call).
//
// The reference reduces gradients with DDP's bucketed NCCL all-reduce (controller.py:78-86; SURVEY.md
// N4) and then runs multiply_grads, clip_grad_norm and Adam as per-tensor loops on EVERY rank over ALL
// parameters (optim.py:59-70,162-231; K15-K17). Here W ranks of one NVSwitch domain share their flat
// gradient / parameter buffers (symmetric memory: every rank maps every peer's copy, and -- when the
// allocation is bound to an NVLS multicast object -- one multicast address that stands for all copies) and
// four kernels do the whole job on each rank:
//
//   hs_dp_reduce_scatter     per bucket, overlapped with backward: announce "my bucket is final" in every
//                            peer's signal pad, wait for all peers, then reduce each OWNED 2048-element chunk
//                            across the W copies -- one multimem.ld_reduce per vector (the switch adds the W
//                            copies, NVLS), or W-1 peer loads when there is no multicast mapping -- keep the
//                            fp32 sum locally and fold its square into the norm partial.
//   hs_dp_norm_exchange      2 W floats cross the fabric: every rank deposits its norm partial and its sample
//                            size in every pad; all ranks form the same global norm, gradient denominator,
//                            clip coefficient and overflow flag. No host round trip, no NCCL.
//   hs_dp_update_allgather   fused unscale * clip * Adam (or Adadelta) on the owned chunks (1/W of the state
//                            and of the HBM traffic); the new 16-bit parameters go to ALL ranks' buffers (one
//                            multimem.st, or W peer stores) -- the all-gather is the update's epilogue. Chunks
//                            are processed in flat (= forward) order and every slice of the buffer ("gather
//                            bucket") is published with its own flag as soon as it is complete.
//   hs_dp_wait_params        returns once the given gather buckets of every peer have landed; the forward pass
//                            queues one in front of the first kernel that reads a slice, so the all-gather of
//                            the later layers overlaps the forward of the earlier ones.
//
// Ownership is cyclic by chunk (chunk c belongs to rank c % W), so any bucket spreads evenly over all
// GPUs' NVLink ports. Cross-GPU ordering: release/acquire at .sys scope on epoch-numbered flags (never
// reset), data accesses at .sys scope; every spin is bounded (hs_dp_set_timeout) and traps instead of
// hanging the GPU if a peer disappears.
//
// The per-rank bodies are device functions that take (rank, block id, block count) as arguments: the
// production kernels pass blockIdx / gridDim, and hs_dp_emulate_step runs W virtual ranks of the SAME
// bodies inside one cooperative launch on one GPU (the 1-GPU protocol test; separate spinning launches on
// one GPU are not guaranteed to be co-scheduled).
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

#include <cooperative_groups.h>
#include <cstring>

namespace hs {

constexpr int DP_THREADS = 128;   // update kernel: 128 threads x <= 64 registers fits next to a GEMM CTA (8 K of the ~11.7 K registers it leaves)
// The reduce-scatter runs WHILE backward's GEMMs own the SMs (320 threads x 168 registers, ~225 KB smem per
// SM). It only overlaps if its CTAs fit into what is left of an SM: no shared memory to speak of and
// <= 11.7 K registers. 256 threads at <= 42 registers (launch bound 6 CTAs/SM) is 10.5 K.
constexpr int RS_THREADS = 256;
constexpr int DP_MAX_WORLD = 16;
constexpr int DP_MAX_GATHER = 64;   // gather buckets (parameter-arrival flags) per rank

// signal pad layout (32-bit words)
__host__ __device__ inline int sig_norm_flag(int W, int src) { return src; }
__host__ __device__ inline int sig_norm_value(int W, int src) { return W + src; }
__host__ __device__ inline int sig_ss_value(int W, int src) { return 2 * W + src; }
__host__ __device__ inline int sig_param_ready(int W, int j, int src) { return 3 * W + j * W + src; }
__host__ __device__ inline int sig_grad_ready(int W, int slot, int src) { return (3 + DP_MAX_GATHER) * W + slot * W + src; }

struct PeerTable {
  void* ptr[DP_MAX_WORLD];
};

// Upper bound of every cross-GPU spin, in SM cycles. The default (~30 min at 2 GHz) is the process-group
// timeout of torch.distributed, which is what bounds the reference's waits inside NCCL; the framework sets
// it from --distributed-timeout. A peer that is merely slow (rank 0 writing a checkpoint) must not trap.
__device__ long long g_dp_timeout_cycles = 3600000000000LL;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
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
// NVLS: `p` is an address inside the multicast mapping of the gradient buffer. The switch fetches the
// vector from every GPU of the group, adds the copies in fp32 and returns ONE response: one request per
// vector instead of W-1, and 1/W of the ingress bytes.
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
// NVLS store: one request leaves this GPU, the switch replicates it into every GPU's copy (ours too).
__device__ __forceinline__ void mc_st_u2(uint2* p, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};"
               ::"l"(p), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y))
               : "memory");
}

// Spin until *p >= want (epochs only grow). Bounded: a vanished peer becomes a trap, not a hung GPU.
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t want, const char* what, int peer) {
  if (static_cast<int32_t>(ld_acquire_sys(p) - want) >= 0) return;
  const long long t0 = clock64();
  const long long limit = g_dp_timeout_cycles;
  while (static_cast<int32_t>(ld_acquire_sys(p) - want) < 0) {
    __nanosleep(64);
    if (clock64() - t0 > limit) {
      printf("hetseq_b200: timed out waiting for %s from rank %d (want epoch %u)\n", what, peer, want);
      __trap();
    }
  }
}

__device__ __forceinline__ float block_sum_to_thread0(float v, float* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();   // `red` may still be read by a previous call
  if (lane == 0) red[warp] = v;
  __syncthreads();
  if (warp == 0) {
    v = lane < (blockDim.x >> 5) ? red[lane] : 0.f;
    v = warp_sum(v);
  }
  return v;
}

// ------------------------------------------------------------------------------------------------
// reduce-scatter of one bucket = chunks [chunk_begin, chunk_end); returns this thread's share of the
// sum of squares of the reduced values
struct RsArgs {
  PeerTable grads, sigs;
  const float* grads_mc;   // multicast address of the gradient buffer, or nullptr (peer loads)
  int W, chunk, slot;
  long long chunk_begin, chunk_end;
};

template <bool MC>
__device__ __forceinline__ float rs_body(const RsArgs& a, int rank, int bid, int nblk, uint32_t epoch) {
  const int W = a.W;
  // 1. announce + wait (block 0 announces; every block waits on its own before touching peer data)
  if (bid == 0 && threadIdx.x < W) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(a.sigs.ptr[threadIdx.x]) + sig_grad_ready(W, a.slot, rank), epoch);
  }
  // our own gradients are final by stream order; only the peers' announcements need waiting for
  if (threadIdx.x < W && threadIdx.x != rank) {
    wait_flag(reinterpret_cast<const uint32_t*>(a.sigs.ptr[rank]) + sig_grad_ready(W, a.slot, threadIdx.x), epoch,
              "gradient bucket", threadIdx.x);
  }
  __syncthreads();

  // 2. owned chunks of this bucket: c in [chunk_begin, chunk_end), c % W == rank
  const long long first = a.chunk_begin + ((rank - a.chunk_begin % W) + W) % W;
  const int vec_per_chunk = a.chunk / 4;
  float acc = 0.f;
  float* mine = reinterpret_cast<float*>(a.grads.ptr[rank]);
  const uint64_t pol = l2_evict_first_policy();
  for (long long c = first + static_cast<long long>(bid) * W; c < a.chunk_end; c += static_cast<long long>(nblk) * W) {
    const long long base = c * a.chunk;
    if constexpr (MC) {
      // four switch-side reductions in flight per thread
      const float4* src = reinterpret_cast<const float4*>(a.grads_mc + base);
      for (int i = threadIdx.x; i < vec_per_chunk; i += 4 * blockDim.x) {
        float4 s[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int iu = i + u * blockDim.x;
          s[u] = iu < vec_per_chunk ? mc_ld_reduce_f4(src + iu) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int iu = i + u * blockDim.x;
          if (iu < vec_per_chunk) {
            st_local_stream_f4(reinterpret_cast<float4*>(mine + base) + iu, s[u], pol);
            acc += s[u].x * s[u].x + s[u].y * s[u].y + s[u].z * s[u].z + s[u].w * s[u].w;
          }
        }
      }
    } else {
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
          const float4* src = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(a.grads.ptr[peer]) + base);
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
  }
  return acc;
}

template <bool MC>
__global__ void __launch_bounds__(RS_THREADS, 6)
dp_reduce_scatter_kernel(RsArgs a, int rank, uint32_t epoch, const uint32_t* __restrict__ epoch_dev,
                         float* __restrict__ sumsq_accum) {
  __shared__ float red[32];
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  float acc = rs_body<MC>(a, rank, blockIdx.x, gridDim.x, epoch);
  acc = block_sum_to_thread0(acc, red);
  if (threadIdx.x == 0) atomicAdd(sumsq_accum, acc);
}

// ------------------------------------------------------------------------------------------------
// out = [grad_norm, multiplier (incl. 1/sample_size and clip), skip flag, global sum of squares,
//        global sample size]; executed by ONE warp per rank
struct NormArgs {
  PeerTable sigs;
  int W;
  float mult_host, max_norm;
  const float* mult_dev;
};

__device__ __forceinline__ void norm_body(const NormArgs& a, int rank, uint32_t epoch, float* sumsq_accum,
                                          const float* sample_size, float* out, int* skip_out, int* pending) {
  const int t = threadIdx.x & 31, W = a.W;
  const float mine = *sumsq_accum;
  const float ss_mine = sample_size != nullptr ? *sample_size : 0.f;
  if (t < W) {
    uint32_t* pad = reinterpret_cast<uint32_t*>(a.sigs.ptr[t]);
    pad[sig_norm_value(W, rank)] = __float_as_uint(mine);
    pad[sig_ss_value(W, rank)] = __float_as_uint(ss_mine);
    __threadfence_system();
    st_release_sys(pad + sig_norm_flag(W, rank), epoch);
  }
  float v = 0.f, s = 0.f;
  if (t < W) {
    const uint32_t* pad = reinterpret_cast<const uint32_t*>(a.sigs.ptr[rank]);
    wait_flag(pad + sig_norm_flag(W, t), epoch, "norm partial", t);
    v = __uint_as_float(*reinterpret_cast<const volatile uint32_t*>(pad + sig_norm_value(W, t)));
    s = __uint_as_float(*reinterpret_cast<const volatile uint32_t*>(pad + sig_ss_value(W, t)));
  }
  // same summation order on every rank -> bit-identical norm and denominator everywhere
  float total = 0.f, ss = 0.f;
  for (int p = 0; p < W; ++p) {
    total += __shfl_sync(0xffffffffu, v, p);
    ss += __shfl_sync(0xffffffffu, s, p);
  }
  if (t == 0) {
    float mult = a.mult_host * (a.mult_dev != nullptr ? *a.mult_dev : 1.0f);
    // gradients are sums over ranks here; the reference's multiply_grads(world / sample_size) on DDP's mean
    // (controller.py:338-339) is 1 / (sum of sample sizes) on the sum
    if (sample_size != nullptr) mult *= ss > 0.f ? 1.0f / ss : 1.0f / static_cast<float>(W);
    const float norm = sqrtf(total) * fabsf(mult);
    const bool bad = !(norm == norm) || isinf(norm);
    float m = mult;
    if (!bad && a.max_norm > 0.f) m *= fminf(a.max_norm / (norm + 1e-6f), 1.0f);
    out[0] = norm;
    out[1] = m;
    out[2] = bad ? 1.f : 0.f;
    out[3] = total;
    out[4] = ss;
    *skip_out = bad ? 1 : 0;
    *sumsq_accum = 0.f;   // ready for the next step's accumulation
    if (pending != nullptr) *pending = 1;   // an update is now waiting to be applied
  }
}

__global__ void __launch_bounds__(32)
dp_norm_exchange_kernel(NormArgs a, int rank, uint32_t epoch, const uint32_t* __restrict__ epoch_dev,
                        float* __restrict__ sumsq_accum, const float* __restrict__ sample_size,
                        float* __restrict__ out, int* __restrict__ skip_out, int* __restrict__ pending) {
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  norm_body(a, rank, epoch, sumsq_accum, sample_size, out, skip_out, pending);
}

// ------------------------------------------------------------------------------------------------
struct DpHyper {
  float lr, a, b, eps, weight_decay, step_factor;   // Adam: a, b = beta1, beta2; Adadelta: a = rho
};
enum { DP_OPT_ADAM = 0, DP_OPT_ADADELTA = 1 };

struct UpdArgs {
  const float* grad;
  float *master, *s1, *s2;        // compact fp32: master weights, (exp_avg, exp_avg_sq) | (square_avg, acc_delta)
  PeerTable params, sigs;
  void* params_mc;                // multicast address of the parameter buffer, or nullptr (peer stores)
  int W, chunk, ngather;
  long long nblocks, blocks_per_gather;
  DpHyper h;
  const float* scalars;           // out of the norm exchange
  const float* hyper_dev;         // [lr, step_factor] or nullptr
  unsigned int* counters;         // [ngather + 1], zero between launches
  int* pending;                   // nullptr: always apply; else apply only when *pending != 0 (and clear it)
};

template <int OPT>
__device__ __forceinline__ void update4(const DpHyper& h, float lr, float step_size, float decay, const float (&gs)[4],
                                        float (&ps)[4], float (&s1)[4], float (&s2)[4]) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if constexpr (OPT == DP_OPT_ADAM) {
      s1[k] = h.a * s1[k] + (1.0f - h.a) * gs[k];
      s2[k] = h.b * s2[k] + (1.0f - h.b) * gs[k] * gs[k];
      ps[k] = ps[k] * decay - step_size * (s1[k] / (sqrtf(s2[k]) + h.eps));   // optim.py:217-226
    } else {
      const float gk = gs[k] + h.weight_decay * ps[k];                         // optim.py:295-296
      s1[k] = h.a * s1[k] + (1.0f - h.a) * gk * gk;
      const float delta = sqrtf(s2[k] + h.eps) / sqrtf(s1[k] + h.eps) * gk;
      ps[k] -= lr * delta;
      s2[k] = h.a * s2[k] + (1.0f - h.a) * delta * delta;
    }
  }
}

// Publish gather bucket j of `rank` to every peer (executed by the CTA that completed it).
__device__ __forceinline__ void publish_gather(const UpdArgs& a, int rank, int j, uint32_t epoch) {
  if (threadIdx.x < a.W) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(a.sigs.ptr[threadIdx.x]) + sig_param_ready(a.W, j, rank), epoch);
  }
}

// A CTA hands in the `mine` blocks of gather bucket `cur` it has finished; whoever completes the bucket publishes it.
__device__ __forceinline__ void hand_in_bucket(const UpdArgs& a, int rank, int cur, unsigned int mine, uint32_t epoch,
                                               int* sh_flag) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long lo = static_cast<long long>(cur) * a.blocks_per_gather;
    const long long hi = lo + a.blocks_per_gather < a.nblocks ? lo + a.blocks_per_gather : a.nblocks;
    const unsigned int prev = atomicAdd(a.counters + cur, mine);
    const bool done = prev + mine == static_cast<unsigned int>(hi - lo);
    if (done) a.counters[cur] = 0;
    *sh_flag = done ? 1 : 0;
  }
  __syncthreads();
  if (*sh_flag) publish_gather(a, rank, cur, epoch);
}

// last CTA out clears "an update is waiting"
__device__ __forceinline__ void clear_pending(const UpdArgs& a, int nblk) {
  if (a.pending == nullptr) return;
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(a.counters + a.ngather, 1u);
    if (prev == static_cast<unsigned int>(nblk) - 1) {
      a.counters[a.ngather] = 0;
      *a.pending = 0;
    }
  }
}

// Update WITHOUT the gather: Adam / Adadelta on the owned chunks at full speed on the main stream, the new 16-bit
// values go into this rank's own parameter buffer only. The peers get them from gather_body, which runs next to
// the following forward pass.
template <typename OutT, int OPT>
__device__ __forceinline__ void update_local_body(const UpdArgs& a, int rank, int bid, int nblk) {
  if (a.scalars[2] != 0.f) return;        // overflow: leave everything untouched
  const int W = a.W;
  const float mult = a.scalars[1];
  const float lr = a.hyper_dev != nullptr ? a.hyper_dev[0] : a.h.lr;
  const float step_size = lr * (a.hyper_dev != nullptr ? a.hyper_dev[1] : a.h.step_factor);
  const float decay = 1.0f - a.h.weight_decay * lr;
  const int vec_per_chunk = a.chunk / 4;
  const long long stride = static_cast<long long>(W) * a.chunk;
  OutT* mine_p = reinterpret_cast<OutT*>(a.params.ptr[rank]);
  for (long long c = bid; c < a.nblocks; c += nblk) {
    const long long goff = c * stride + static_cast<long long>(rank) * a.chunk;
    const float4* g4 = reinterpret_cast<const float4*>(a.grad + goff);
    float4* p4 = reinterpret_cast<float4*>(a.master + c * static_cast<long long>(a.chunk));
    float4* m4 = reinterpret_cast<float4*>(a.s1 + c * static_cast<long long>(a.chunk));
    float4* v4 = reinterpret_cast<float4*>(a.s2 + c * static_cast<long long>(a.chunk));
    for (int i = threadIdx.x; i < vec_per_chunk; i += blockDim.x) {
      const float4 g = g4[i];
      const float4 mm = m4[i], vv = v4[i], pp = p4[i];
      const float gs[4] = {g.x * mult, g.y * mult, g.z * mult, g.w * mult};
      float s1[4] = {mm.x, mm.y, mm.z, mm.w}, s2[4] = {vv.x, vv.y, vv.z, vv.w}, ps[4] = {pp.x, pp.y, pp.z, pp.w};
      update4<OPT>(a.h, lr, step_size, decay, gs, ps, s1, s2);
      m4[i] = make_float4(s1[0], s1[1], s1[2], s1[3]);
      v4[i] = make_float4(s2[0], s2[1], s2[2], s2[3]);
      p4[i] = make_float4(ps[0], ps[1], ps[2], ps[3]);
      uint2 o;
      o.x = Cvt<OutT>::pack2(ps[0], ps[1]);
      o.y = Cvt<OutT>::pack2(ps[2], ps[3]);
      reinterpret_cast<uint2*>(mine_p + goff)[i] = o;
    }
  }
}

// The all-gather on its own: copy this rank's owned 16-bit chunks into every peer's parameter buffer (one
// multimem.st per vector, or W-1 peer stores), in flat (= forward) order, publishing each gather bucket as soon
// as it is complete. A pure copy: ~20 registers and no shared memory, so its CTAs fit next to the forward
// pass's GEMM CTAs, and NVLink -- not the SMs -- sets its pace.
template <bool MC>
__device__ __forceinline__ void gather_body(const UpdArgs& a, int rank, int bid, int nblk, uint32_t epoch, int* sh_flag) {
  const int W = a.W;
  const bool apply = (a.pending == nullptr || *a.pending != 0) && a.scalars[2] == 0.f;
  if (!apply) {
    if (bid == 0)
      for (int j = 0; j < a.ngather; ++j) publish_gather(a, rank, j, epoch);
  } else {
    const int vec_per_chunk = a.chunk / 8;      // uint4 = 8 16-bit parameters
    const long long stride = static_cast<long long>(W) * a.chunk;
    const uint16_t* mine_p = reinterpret_cast<const uint16_t*>(a.params.ptr[rank]);
    int cur = -1;
    unsigned int mine = 0;
    for (long long c = bid;; c += nblk) {
      const int j = c < a.nblocks ? static_cast<int>(c / a.blocks_per_gather) : -2;
      if (j != cur) {
        if (cur >= 0) hand_in_bucket(a, rank, cur, mine, epoch, sh_flag);
        cur = j;
        mine = 0;
      }
      if (j < 0) break;
      const long long goff = c * stride + static_cast<long long>(rank) * a.chunk;
      const uint4* src = reinterpret_cast<const uint4*>(mine_p + goff);
      for (int i = threadIdx.x; i < vec_per_chunk; i += blockDim.x) {
        const uint4 v = src[i];
        if constexpr (MC) {
          asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                       ::"l"(reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(a.params_mc) + goff) + i),
                         "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                         "f"(__uint_as_float(v.w))
                       : "memory");
        } else {
#pragma unroll 4
          for (int p = 1; p < W; ++p) {
            const int peer = (rank + p) % W;
            asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};"
                         ::"l"(reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(a.params.ptr[peer]) + goff) + i),
                           "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                         : "memory");
          }
        }
      }
      ++mine;
    }
  }
  clear_pending(a, nblk);
}

template <typename OutT, int OPT, bool MC>
__device__ __forceinline__ void update_body(const UpdArgs& a, int rank, int bid, int nblk, uint32_t epoch, int* sh_flag) {
  const int W = a.W;
  const bool apply = (a.pending == nullptr || *a.pending != 0) && a.scalars[2] == 0.f;
  if (!apply) {
    // overflow step, or nothing waiting: parameters are what they were; say so for every gather bucket
    if (bid == 0)
      for (int j = 0; j < a.ngather; ++j) publish_gather(a, rank, j, epoch);
  } else {
    const float mult = a.scalars[1];
    const float lr = a.hyper_dev != nullptr ? a.hyper_dev[0] : a.h.lr;
    const float step_size = lr * (a.hyper_dev != nullptr ? a.hyper_dev[1] : a.h.step_factor);
    const float decay = 1.0f - a.h.weight_decay * lr;
    const int vec_per_chunk = a.chunk / 4;
    const long long stride = static_cast<long long>(W) * a.chunk;
    int cur = -1;            // gather bucket this CTA is working in
    unsigned int mine = 0;   // blocks of `cur` this CTA has finished
    for (long long c = bid;; c += nblk) {
      const int j = c < a.nblocks ? static_cast<int>(c / a.blocks_per_gather) : -2;
      if (j != cur) {
        if (cur >= 0) hand_in_bucket(a, rank, cur, mine, epoch, sh_flag);
        cur = j;
        mine = 0;
      }
      if (j < 0) break;
      const long long goff = c * stride + static_cast<long long>(rank) * a.chunk;   // position in the flat buffers
      const float4* g4 = reinterpret_cast<const float4*>(a.grad + goff);
      float4* p4 = reinterpret_cast<float4*>(a.master + c * static_cast<long long>(a.chunk));
      float4* m4 = reinterpret_cast<float4*>(a.s1 + c * static_cast<long long>(a.chunk));
      float4* v4 = reinterpret_cast<float4*>(a.s2 + c * static_cast<long long>(a.chunk));
      for (int i = threadIdx.x; i < vec_per_chunk; i += blockDim.x) {
        const float4 g = g4[i];
        const float4 mm = m4[i], vv = v4[i], pp = p4[i];
        const float gs[4] = {g.x * mult, g.y * mult, g.z * mult, g.w * mult};
        float s1[4] = {mm.x, mm.y, mm.z, mm.w}, s2[4] = {vv.x, vv.y, vv.z, vv.w}, ps[4] = {pp.x, pp.y, pp.z, pp.w};
        update4<OPT>(a.h, lr, step_size, decay, gs, ps, s1, s2);
        m4[i] = make_float4(s1[0], s1[1], s1[2], s1[3]);
        v4[i] = make_float4(s2[0], s2[1], s2[2], s2[3]);
        p4[i] = make_float4(ps[0], ps[1], ps[2], ps[3]);
        uint2 o;
        o.x = Cvt<OutT>::pack2(ps[0], ps[1]);
        o.y = Cvt<OutT>::pack2(ps[2], ps[3]);
        // the all-gather: the update's epilogue stores the new parameters into every rank's buffer
        if constexpr (MC) {
          mc_st_u2(reinterpret_cast<uint2*>(reinterpret_cast<OutT*>(a.params_mc) + goff) + i, o);
        } else {
#pragma unroll 4
          for (int p = 0; p < W; ++p) {
            const int peer = (rank + p) % W;
            st_relaxed_sys_u2(reinterpret_cast<uint2*>(reinterpret_cast<OutT*>(a.params.ptr[peer]) + goff) + i, o);
          }
        }
      }
      ++mine;
    }
  }
  clear_pending(a, nblk);
}

template <typename OutT, int OPT, bool MC>
__global__ void __launch_bounds__(DP_THREADS, 8)
dp_update_allgather_kernel(UpdArgs a, int rank, uint32_t epoch, const uint32_t* __restrict__ epoch_dev, int epoch_offset) {
  __shared__ int sh_flag;
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  update_body<OutT, OPT, MC>(a, rank, blockIdx.x, gridDim.x, epoch + epoch_offset, &sh_flag);
}

template <typename OutT, int OPT>
__global__ void __launch_bounds__(256)
dp_update_local_kernel(UpdArgs a, int rank) {
  update_local_body<OutT, OPT>(a, rank, blockIdx.x, gridDim.x);
}

template <bool MC>
__global__ void __launch_bounds__(RS_THREADS, 6)
dp_allgather_kernel(UpdArgs a, int rank, uint32_t epoch, const uint32_t* __restrict__ epoch_dev, int epoch_offset) {
  __shared__ int sh_flag;
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  gather_body<MC>(a, rank, blockIdx.x, gridDim.x, epoch + epoch_offset, &sh_flag);
}

// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void wait_params_body(const PeerTable& sigs, int rank, int W, int j_lo, int j_hi, uint32_t epoch) {
  const int n = (j_hi - j_lo + 1) * W;
  for (int idx = threadIdx.x; idx < n; idx += blockDim.x) {
    const int j = j_lo + idx / W, src = idx % W;
    wait_flag(reinterpret_cast<const uint32_t*>(sigs.ptr[rank]) + sig_param_ready(W, j, src), epoch,
              "updated parameters", src);
  }
}

__global__ void __launch_bounds__(128)
dp_wait_params_kernel(PeerTable sigs, int rank, int W, int j_lo, int j_hi, uint32_t epoch,
                      const uint32_t* __restrict__ epoch_dev, int epoch_offset) {
  if (epoch_dev != nullptr) epoch = *epoch_dev;
  wait_params_body(sigs, rank, W, j_lo, j_hi, epoch + epoch_offset);
}

// ------------------------------------------------------------------------------------------------
// W virtual ranks in one cooperative launch (all CTAs co-resident, so the cross-"rank" spins are safe).
// CTAs [r*G, (r+1)*G) play rank r and run the production bodies above; kernel boundaries of the real
// sequence (reduce-scatter launches | norm exchange | update) become barriers among the G CTAs of a rank.
struct EmuRank {
  float* sumsq;
  const float* sample_size;
  float* scalars;
  int* skip;
  float *master, *s1, *s2;
  unsigned int* counters;
  int* pending;
};
struct EmuArgs {
  EmuRank r[DP_MAX_WORLD];
  PeerTable grads, params, sigs;
  int W, G, chunk, nbuckets, ngather, opt, split;
  long long total_chunks, chunks_per_bucket, nblocks, blocks_per_gather;
  NormArgs norm;
  DpHyper h;
  const float* hyper_dev;
  unsigned int* barrier_words;   // [W], zero at launch
  uint32_t epoch;
};

__device__ __forceinline__ void rank_barrier(unsigned int* word, unsigned int target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(word, 1u);
    while (*reinterpret_cast<volatile unsigned int*>(word) < target) __nanosleep(32);
    __threadfence();
  }
  __syncthreads();
}

template <typename OutT>
__global__ void __launch_bounds__(RS_THREADS)
dp_emulate_kernel(EmuArgs e) {
  __shared__ float red[32];
  __shared__ int sh_flag;
  const int rank = blockIdx.x / e.G, bid = blockIdx.x % e.G;
  const EmuRank& me = e.r[rank];
  unsigned int* bar = e.barrier_words + rank;
  // reduce-scatter, bucket by bucket from the END of the buffer (the order backward produces them)
  for (int b = 0; b < e.nbuckets; ++b) {
    RsArgs a;
    a.grads = e.grads; a.sigs = e.sigs; a.grads_mc = nullptr; a.W = e.W; a.chunk = e.chunk; a.slot = b;
    a.chunk_end = e.total_chunks - static_cast<long long>(b) * e.chunks_per_bucket;
    a.chunk_begin = a.chunk_end - e.chunks_per_bucket > 0 ? a.chunk_end - e.chunks_per_bucket : 0;
    float acc = rs_body<false>(a, rank, bid, e.G, e.epoch);
    acc = block_sum_to_thread0(acc, red);
    if (threadIdx.x == 0) atomicAdd(me.sumsq, acc);
  }
  rank_barrier(bar, e.G);
  if (bid == 0 && threadIdx.x < 32)
    norm_body(e.norm, rank, e.epoch, me.sumsq, me.sample_size, me.scalars, me.skip, me.pending);
  rank_barrier(bar, 2 * e.G);
  UpdArgs u;
  u.grad = reinterpret_cast<const float*>(e.grads.ptr[rank]);
  u.master = me.master; u.s1 = me.s1; u.s2 = me.s2;
  u.params = e.params; u.sigs = e.sigs; u.params_mc = nullptr;
  u.W = e.W; u.chunk = e.chunk; u.ngather = e.ngather; u.nblocks = e.nblocks; u.blocks_per_gather = e.blocks_per_gather;
  u.h = e.h; u.scalars = me.scalars; u.hyper_dev = e.hyper_dev; u.counters = me.counters; u.pending = me.pending;
  if (e.split) {
    // the production sequence: update at the end of the step, gather next to the following forward pass
    if (e.opt == DP_OPT_ADAM) update_local_body<OutT, DP_OPT_ADAM>(u, rank, bid, e.G);
    else update_local_body<OutT, DP_OPT_ADADELTA>(u, rank, bid, e.G);
    rank_barrier(bar, 3 * e.G);
    gather_body<false>(u, rank, bid, e.G, e.epoch, &sh_flag);
  } else if (e.opt == DP_OPT_ADAM) {
    update_body<OutT, DP_OPT_ADAM, false>(u, rank, bid, e.G, e.epoch, &sh_flag);
  } else {
    update_body<OutT, DP_OPT_ADADELTA, false>(u, rank, bid, e.G, e.epoch, &sh_flag);
  }
  if (bid == 0) wait_params_body(e.sigs, rank, e.W, 0, e.ngather - 1, e.epoch);
}

static int fill_table(PeerTable* t, void* const* ptrs, int W) {
  if (W < 1 || W > DP_MAX_WORLD) { set_last_error("fused data-parallel path supports 1..%d ranks (got %d)", DP_MAX_WORLD, W); return 1; }
  for (int i = 0; i < DP_MAX_WORLD; ++i) t->ptr[i] = i < W ? ptrs[i] : nullptr;
  return 0;
}

static long long gather_blocks(long long nblocks, int ngather) { return (nblocks + ngather - 1) / ngather; }
// number of NON-EMPTY gather buckets: with blocks_per_gather rounded up the last requested buckets may own
// no block at all, and a bucket nobody works on is never published
static int gather_count(long long nblocks, int ngather) {
  const long long bpg = gather_blocks(nblocks, ngather);
  return static_cast<int>((nblocks + bpg - 1) / bpg);
}

}  // namespace hs

using namespace hs;

// ---- symmetric memory plumbing: CUDA IPC fallback (no multicast) -----------------------------------
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
extern "C" int hs_dp_signal_words(int W, int max_buckets) { return (3 + DP_MAX_GATHER) * W + max_buckets * W; }
extern "C" int hs_dp_max_gather_buckets(void) { return DP_MAX_GATHER; }

extern "C" int hs_dp_set_timeout(double seconds) {
  int dev = 0, khz = 0;
  HS_CUDA_OK(cudaGetDevice(&dev));
  HS_CUDA_OK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  if (khz <= 0) khz = 2000000;
  if (seconds < 1.0) seconds = 1.0;
  const long long cycles = static_cast<long long>(seconds * 1000.0 * static_cast<double>(khz));
  HS_CUDA_OK(cudaMemcpyToSymbol(g_dp_timeout_cycles, &cycles, sizeof(cycles)));
  return 0;
}

// ---- kernels ----------------------------------------------------------------------------------------
extern "C" int hs_dp_reduce_scatter(void* const* grad_ptrs, const void* grads_mc, void* const* sig_ptrs, int rank, int W,
                                    long long chunk_begin, long long chunk_end, int chunk, int slot, unsigned epoch,
                                    const unsigned* epoch_dev, float* sumsq_accum, int nblocks, hs_stream_t stream) {
  RsArgs a;
  if (fill_table(&a.grads, grad_ptrs, W) || fill_table(&a.sigs, sig_ptrs, W)) return 1;
  if (chunk % 4 != 0 || chunk_end <= chunk_begin || rank < 0 || rank >= W) {
    set_last_error("hs_dp_reduce_scatter: bad arguments rank=%d W=%d chunks=[%lld,%lld) chunk=%d", rank, W, chunk_begin, chunk_end, chunk);
    return 1;
  }
  a.grads_mc = static_cast<const float*>(grads_mc);
  a.W = W; a.chunk = chunk; a.slot = slot; a.chunk_begin = chunk_begin; a.chunk_end = chunk_end;
  // nblocks <= 0: one CTA per owned chunk (short-lived CTAs). What decides how well this kernel overlaps with
  // backward turned out to be stream PRIORITY, not CTA lifetime: launched at a higher priority than the compute
  // stream its pending CTAs took every SM slot freed at a GEMM kernel boundary and the next GEMM's CTAs (which
  // need almost a whole SM each) waited; at a lower priority it only uses what backward leaves free.
  const long long owned = (chunk_end - chunk_begin + W - 1) / W;
  long long want = nblocks > 0 ? nblocks : owned;
  if (want > owned) want = owned;
  if (want > 1 << 20) want = 1 << 20;
  const int grid = static_cast<int>(want < 1 ? 1 : want);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (grads_mc != nullptr) dp_reduce_scatter_kernel<true><<<grid, RS_THREADS, 0, st>>>(a, rank, epoch, epoch_dev, sumsq_accum);
  else dp_reduce_scatter_kernel<false><<<grid, RS_THREADS, 0, st>>>(a, rank, epoch, epoch_dev, sumsq_accum);
  return post_launch("dp_reduce_scatter_kernel");
}

extern "C" int hs_dp_norm_exchange(void* const* sig_ptrs, int rank, int W, unsigned epoch, const unsigned* epoch_dev,
                                   float* sumsq_accum, float mult_host, const float* mult_dev, const float* sample_size,
                                   float max_norm, float* out5, int* skip_out, int* pending, hs_stream_t stream) {
  NormArgs a;
  if (fill_table(&a.sigs, sig_ptrs, W)) return 1;
  a.W = W; a.mult_host = mult_host; a.mult_dev = mult_dev; a.max_norm = max_norm;
  dp_norm_exchange_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a, rank, epoch, epoch_dev, sumsq_accum, sample_size, out5, skip_out, pending);
  return post_launch("dp_norm_exchange_kernel");
}

template <typename OutT, bool MC>
static void launch_update(int opt, int grid, cudaStream_t st, const UpdArgs& a, int rank, unsigned epoch,
                          const unsigned* epoch_dev, int epoch_offset) {
  if (opt == DP_OPT_ADAM) dp_update_allgather_kernel<OutT, DP_OPT_ADAM, MC><<<grid, DP_THREADS, 0, st>>>(a, rank, epoch, epoch_dev, epoch_offset);
  else dp_update_allgather_kernel<OutT, DP_OPT_ADADELTA, MC><<<grid, DP_THREADS, 0, st>>>(a, rank, epoch, epoch_dev, epoch_offset);
}

extern "C" int hs_dp_update_allgather(int opt, const float* grad, float* master, float* state1, float* state2,
                                      void* const* param_ptrs, void* params_mc, void* const* sig_ptrs, int rank, int W,
                                      long long nblocks, int chunk, int out_dtype, float lr, float a_coef, float b_coef,
                                      float eps, float weight_decay, float step_factor, const float* scalars,
                                      const float* hyper_dev, unsigned epoch, const unsigned* epoch_dev, int epoch_offset,
                                      unsigned int* counters, int* pending, int ngather, int grid_blocks,
                                      hs_stream_t stream) {
  UpdArgs a;
  if (fill_table(&a.params, param_ptrs, W) || fill_table(&a.sigs, sig_ptrs, W)) return 1;
  if (chunk % 4 != 0 || nblocks <= 0 || ngather < 1 || ngather > DP_MAX_GATHER || (opt != DP_OPT_ADAM && opt != DP_OPT_ADADELTA)) {
    set_last_error("hs_dp_update_allgather: bad layout (chunk=%d nblocks=%lld ngather=%d opt=%d)", chunk, nblocks, ngather, opt);
    return 1;
  }
  a.grad = grad; a.master = master; a.s1 = state1; a.s2 = state2; a.params_mc = params_mc;
  a.W = W; a.chunk = chunk; a.ngather = gather_count(nblocks, ngather); a.nblocks = nblocks;
  a.blocks_per_gather = gather_blocks(nblocks, ngather);
  a.h = DpHyper{lr, a_coef, b_coef, eps, weight_decay, step_factor};
  a.scalars = scalars; a.hyper_dev = hyper_dev; a.counters = counters; a.pending = pending;
  int grid = grid_blocks > 0 ? grid_blocks : num_sms() * 8;
  if (grid > nblocks) grid = static_cast<int>(nblocks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool mc = params_mc != nullptr;
  if (out_dtype == HS_F16) {
    if (mc) launch_update<__half, true>(opt, grid, st, a, rank, epoch, epoch_dev, epoch_offset);
    else launch_update<__half, false>(opt, grid, st, a, rank, epoch, epoch_dev, epoch_offset);
  } else if (out_dtype == HS_BF16) {
    if (mc) launch_update<__nv_bfloat16, true>(opt, grid, st, a, rank, epoch, epoch_dev, epoch_offset);
    else launch_update<__nv_bfloat16, false>(opt, grid, st, a, rank, epoch, epoch_dev, epoch_offset);
  } else { set_last_error("hs_dp_update_allgather: parameters must be f16 or bf16"); return 1; }
  return post_launch("dp_update_allgather_kernel");
}

// Adam / Adadelta on the owned chunks only (no stores to peers, no flags): the end-of-step half of the split update.
extern "C" int hs_dp_update_local(int opt, const float* grad, float* master, float* state1, float* state2,
                                  void* param_local, int rank, int W, long long nblocks, int chunk, int out_dtype,
                                  float lr, float a_coef, float b_coef, float eps, float weight_decay, float step_factor,
                                  const float* scalars, const float* hyper_dev, int grid_blocks, hs_stream_t stream) {
  UpdArgs a;
  memset(&a, 0, sizeof(a));
  if (chunk % 4 != 0 || nblocks <= 0 || rank < 0 || rank >= W || W > DP_MAX_WORLD || (opt != DP_OPT_ADAM && opt != DP_OPT_ADADELTA)) {
    set_last_error("hs_dp_update_local: bad arguments (chunk=%d nblocks=%lld rank=%d W=%d opt=%d)", chunk, nblocks, rank, W, opt);
    return 1;
  }
  a.grad = grad; a.master = master; a.s1 = state1; a.s2 = state2;
  a.params.ptr[rank] = param_local;
  a.W = W; a.chunk = chunk; a.nblocks = nblocks;
  a.h = DpHyper{lr, a_coef, b_coef, eps, weight_decay, step_factor};
  a.scalars = scalars; a.hyper_dev = hyper_dev;
  int grid = grid_blocks > 0 ? grid_blocks : num_sms() * 8;
  if (grid > nblocks) grid = static_cast<int>(nblocks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_dtype == HS_F16) {
    if (opt == DP_OPT_ADAM) dp_update_local_kernel<__half, DP_OPT_ADAM><<<grid, 256, 0, st>>>(a, rank);
    else dp_update_local_kernel<__half, DP_OPT_ADADELTA><<<grid, 256, 0, st>>>(a, rank);
  } else if (out_dtype == HS_BF16) {
    if (opt == DP_OPT_ADAM) dp_update_local_kernel<__nv_bfloat16, DP_OPT_ADAM><<<grid, 256, 0, st>>>(a, rank);
    else dp_update_local_kernel<__nv_bfloat16, DP_OPT_ADADELTA><<<grid, 256, 0, st>>>(a, rank);
  } else { set_last_error("hs_dp_update_local: parameters must be f16 or bf16"); return 1; }
  return post_launch("dp_update_local_kernel");
}

// The gather half: this rank's owned 16-bit chunks go to every peer, one arrival flag per gather bucket.
extern "C" int hs_dp_allgather(void* const* param_ptrs, void* params_mc, void* const* sig_ptrs, int rank, int W,
                               long long nblocks, int chunk, const float* scalars, unsigned epoch,
                               const unsigned* epoch_dev, int epoch_offset, unsigned int* counters, int* pending,
                               int ngather, int grid_blocks, hs_stream_t stream) {
  UpdArgs a;
  memset(&a, 0, sizeof(a));
  if (fill_table(&a.params, param_ptrs, W) || fill_table(&a.sigs, sig_ptrs, W)) return 1;
  if (chunk % 8 != 0 || nblocks <= 0 || ngather < 1 || ngather > DP_MAX_GATHER) {
    set_last_error("hs_dp_allgather: bad layout (chunk=%d nblocks=%lld ngather=%d)", chunk, nblocks, ngather);
    return 1;
  }
  a.params_mc = params_mc;
  a.W = W; a.chunk = chunk; a.ngather = gather_count(nblocks, ngather); a.nblocks = nblocks;
  a.blocks_per_gather = gather_blocks(nblocks, ngather);
  a.scalars = scalars; a.counters = counters; a.pending = pending;
  int grid = grid_blocks > 0 ? grid_blocks : num_sms();
  if (grid > nblocks) grid = static_cast<int>(nblocks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (params_mc != nullptr) dp_allgather_kernel<true><<<grid, RS_THREADS, 0, st>>>(a, rank, epoch, epoch_dev, epoch_offset);
  else dp_allgather_kernel<false><<<grid, RS_THREADS, 0, st>>>(a, rank, epoch, epoch_dev, epoch_offset);
  return post_launch("dp_allgather_kernel");
}

extern "C" int hs_dp_gather_bucket_of(long long nblocks, int ngather, long long block) {
  const long long bpg = gather_blocks(nblocks, ngather);
  const int n = gather_count(nblocks, ngather);
  long long j = block / bpg;
  if (j >= n) j = n - 1;
  return static_cast<int>(j);
}
extern "C" int hs_dp_gather_count(long long nblocks, int ngather) {
  if (nblocks <= 0 || ngather <= 0) return 0;
  return gather_count(nblocks, ngather);
}

extern "C" int hs_dp_wait_params(void* const* sig_ptrs, int rank, int W, int j_lo, int j_hi, unsigned epoch,
                                 const unsigned* epoch_dev, int epoch_offset, hs_stream_t stream) {
  PeerTable s;
  if (fill_table(&s, sig_ptrs, W)) return 1;
  if (j_lo < 0 || j_hi < j_lo || j_hi >= DP_MAX_GATHER) { set_last_error("hs_dp_wait_params: bad bucket range [%d, %d]", j_lo, j_hi); return 1; }
  dp_wait_params_kernel<<<1, 128, 0, static_cast<cudaStream_t>(stream)>>>(s, rank, W, j_lo, j_hi, epoch, epoch_dev, epoch_offset);
  return post_launch("dp_wait_params_kernel");
}

// One whole update of W virtual ranks on ONE GPU (cooperative launch). All pointer arrays are host arrays of
// W device pointers. The buckets are `chunks_per_bucket` chunks each, taken from the end of the buffer.
extern "C" int hs_dp_emulate_step(int W, int blocks_per_rank, void* const* grad_ptrs, void* const* param_ptrs,
                                  void* const* sig_ptrs, float* const* sumsq, const float* const* sample_size,
                                  float* const* scalars, int* const* skip, float* const* master, float* const* state1,
                                  float* const* state2, unsigned int* const* counters, int* const* pending,
                                  unsigned int* barrier_words, long long total_chunks, long long chunks_per_bucket,
                                  int chunk, int ngather, int opt, int out_dtype, float lr, float a_coef, float b_coef,
                                  float eps, float weight_decay, float step_factor, float mult_host, float max_norm,
                                  unsigned epoch, int split, hs_stream_t stream) {
  EmuArgs e;
  memset(&e, 0, sizeof(e));
  if (fill_table(&e.grads, grad_ptrs, W) || fill_table(&e.params, param_ptrs, W) || fill_table(&e.sigs, sig_ptrs, W)) return 1;
  if (blocks_per_rank < 1 || total_chunks % W != 0 || chunks_per_bucket < 1 || ngather < 1 || ngather > DP_MAX_GATHER) {
    set_last_error("hs_dp_emulate_step: bad arguments");
    return 1;
  }
  for (int r = 0; r < W; ++r)
    e.r[r] = EmuRank{sumsq[r], sample_size ? sample_size[r] : nullptr, scalars[r], skip[r], master[r], state1[r], state2[r], counters[r], pending ? pending[r] : nullptr};
  e.W = W; e.G = blocks_per_rank; e.chunk = chunk; e.ngather = ngather; e.opt = opt; e.split = split;
  e.total_chunks = total_chunks; e.chunks_per_bucket = chunks_per_bucket;
  e.nbuckets = static_cast<int>((total_chunks + chunks_per_bucket - 1) / chunks_per_bucket);
  e.nblocks = total_chunks / W; e.blocks_per_gather = gather_blocks(e.nblocks, ngather);
  e.ngather = gather_count(e.nblocks, ngather);
  e.norm.sigs = e.sigs; e.norm.W = W; e.norm.mult_host = mult_host; e.norm.max_norm = max_norm; e.norm.mult_dev = nullptr;
  e.h = DpHyper{lr, a_coef, b_coef, eps, weight_decay, step_factor};
  e.hyper_dev = nullptr; e.barrier_words = barrier_words; e.epoch = epoch;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HS_CUDA_OK(cudaMemsetAsync(barrier_words, 0, sizeof(unsigned int) * W, st));
  void* kargs[] = {&e};
  const void* fn = out_dtype == HS_F16 ? reinterpret_cast<const void*>(dp_emulate_kernel<__half>)
                                       : reinterpret_cast<const void*>(dp_emulate_kernel<__nv_bfloat16>);
  int per_sm = 0;
  HS_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, RS_THREADS, 0));
  if (static_cast<long long>(W) * blocks_per_rank > static_cast<long long>(per_sm) * num_sms()) {
    set_last_error("hs_dp_emulate_step: %d x %d CTAs cannot be co-resident", W, blocks_per_rank);
    return 1;
  }
  HS_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3(W * blocks_per_rank), dim3(RS_THREADS), kargs, 0, st));
  return post_launch("dp_emulate_kernel");
}
