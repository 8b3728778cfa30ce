// hetseq_b200 -- BERT embeddings: gather-sum forward and gradient scatter.
//
// Reference: BertEmbeddings (bert_modeling.py:292-320): word + position(arange) + token-type lookups,
// then LayerNorm and dropout -- 3 embedding kernels + 2 adds forward, and 3 embedding_dense_backward
// launches (atomic scatter-add into [V,H]) backward.
//   hs_embed_gather_sum    z[m,:] = W[ids[m],:] + P[m % S,:] + T[seg[m],:]        (LayerNorm + dropout
//                          follow through hs_ln_fwd(post_dropout=1), the gradient through hs_ln_bwd)
//   hs_embed_backward_sorted  dW[id,:] += sum of de[m,:] over the positions m holding token id, as a segmented
//                          reduction over the positions SORTED by token id, cut into 16-row segments so that the
//                          long [PAD] / [MASK] runs are summed by many CTAs; one writer per slice of dW, fixed
//                          order of additions -- deterministic, no atomics. The same pass over de
//                          produces the two token-type sums dT[t,:]. The order and the run boundaries come from
//                          the host (computed in the batch-prefetch thread) or from a device sort.
//   hs_embed_scatter_word  the same sum with fp32 vector atomics (order of additions varies from run to run)
//   hs_sum_over_batch      dP[s,:] (+)= sum_b de[b*S+s,:]            deterministic, no atomics
//   hs_colsum_rows_where   dT[t,:] (+)= sum_{m: seg[m]==t} de[m,:]   deterministic two-stage column sum (used when
//                          there are more than two token types or no sorted plan)
// Token ids arrive as int32 (the trainer narrows them before the host->device copy) or int64.
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

namespace hs {

template <typename T>
__device__ __forceinline__ void unpack8e(const uint4& u, float (&f)[8]) {
  float2 a = Cvt<T>::unpack2(u.x), b = Cvt<T>::unpack2(u.y), c = Cvt<T>::unpack2(u.z), d = Cvt<T>::unpack2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

template <typename T, typename IdT>
__global__ void __launch_bounds__(256)
embed_gather_sum_kernel(const T* __restrict__ W, const T* __restrict__ P, const T* __restrict__ Tt,
                        const IdT* __restrict__ ids, const IdT* __restrict__ seg, T* __restrict__ z, int M, int S,
                        int H, int vocab, int ntypes) {
  const int vpr = H >> 3;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int m = static_cast<int>(t / vpr), c = static_cast<int>(t % vpr);
  if (m >= M) return;
  long long id = static_cast<long long>(ids[m]);
  long long sg = static_cast<long long>(seg[m]);
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  sg = sg < 0 ? 0 : (sg >= ntypes ? ntypes - 1 : sg);
  float a[8], b[8], d[8];
  unpack8e<T>(reinterpret_cast<const uint4*>(W + static_cast<size_t>(id) * H)[c], a);
  unpack8e<T>(reinterpret_cast<const uint4*>(P + static_cast<size_t>(m % S) * H)[c], b);
  unpack8e<T>(reinterpret_cast<const uint4*>(Tt + static_cast<size_t>(sg) * H)[c], d);
  uint4 o;
  o.x = Cvt<T>::pack2(a[0] + b[0] + d[0], a[1] + b[1] + d[1]);
  o.y = Cvt<T>::pack2(a[2] + b[2] + d[2], a[3] + b[3] + d[3]);
  o.z = Cvt<T>::pack2(a[4] + b[4] + d[4], a[5] + b[5] + d[5]);
  o.w = Cvt<T>::pack2(a[6] + b[6] + d[6], a[7] + b[7] + d[7]);
  reinterpret_cast<uint4*>(z + static_cast<size_t>(m) * H)[c] = o;
}

template <typename T, typename IdT>
__global__ void __launch_bounds__(256)
embed_scatter_word_kernel(const T* __restrict__ de, const IdT* __restrict__ ids, float* __restrict__ dW, int M, int H,
                          int vocab) {
  const int vpr = H >> 3;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int m = static_cast<int>(t / vpr), c = static_cast<int>(t % vpr);
  if (m >= M) return;
  const long long id = static_cast<long long>(ids[m]);
  if (id < 0 || id >= vocab) return;
  float f[8];
  unpack8e<T>(reinterpret_cast<const uint4*>(de + static_cast<size_t>(m) * H)[c], f);
  float4* dst = reinterpret_cast<float4*>(dW + static_cast<size_t>(id) * H + c * 8);
  atomicAdd(dst, make_float4(f[0], f[1], f[2], f[3]));
  atomicAdd(dst + 1, make_float4(f[4], f[5], f[6], f[7]));
}

// ---- deterministic gradient of the word (and token-type) embeddings ------------------------------------------
// order[i]: position (row of de) that is i-th when positions are sorted by token id (stable: ties in ascending
// position). run_end[i]: for the FIRST sorted slot of a run of equal ids, one past its last slot; 0 elsewhere.
//
// Real batches have very long runs ([PAD] is a quarter of all positions, [MASK] another tenth), so a run is not
// the unit of work: the sorted slots are cut into segments of ESEG rows, one CTA per segment, every row read once.
//   phase 1  a thread owns 8 columns, has the segment's ESEG rows in flight at once and adds them in slot order.
//            A run that lies inside the segment goes straight into dW (the CTA is its only writer); the piece of a
//            run that crosses the segment's first or last slot goes to partW[segment][0 = enters from the left,
//            1 = starts here and leaves to the right]; the per-token-type sums of the segment go to partT.
//   phase 2  the segment in which a crossing run STARTS adds that run's pieces in segment order (8 loads in
//            flight) and is the only writer of dW[id]; 64 more CTAs add up the token-type partials, segments
//            dealt round-robin to 8 thread groups whose sums are combined in a fixed order.
// Every sum has one fixed association order, there are no atomics: the result is bit-identical from run to run.
constexpr int ESEG = 16;
constexpr int EMB_THREADS = 128;
constexpr int TYPE_COLS = 16, TYPE_GROUPS = EMB_THREADS / TYPE_COLS;

// ids of the slots around segment s: sid[k] = token id of sorted slot s * ESEG - 1 + k, or -1 when there is no such
// slot, its position is out of range or its id is not a row of dW. pos / sgv: position and token type of the slots.
template <typename IdT>
__device__ __forceinline__ void load_segment_ids(const IdT* __restrict__ ids, const IdT* __restrict__ seg,
                                                 const int* __restrict__ order, int s, int M, int vocab,
                                                 long long* sid, int* pos, int* sgv) {
  const int k = threadIdx.x;
  if (k < ESEG + 2) {
    const int j = s * ESEG - 1 + k;
    long long id = -1;
    int p = -1, sg = -1;
    if (j >= 0 && j < M) {
      p = order[j];
      if (p >= 0 && p < M) {
        id = static_cast<long long>(ids[p]);
        if (id < 0 || id >= vocab) id = -1;
        if (seg != nullptr) sg = static_cast<int>(seg[p]);
      } else {
        p = -1;
      }
    }
    sid[k] = id;
    if (k >= 1 && k <= ESEG && pos != nullptr) { pos[k - 1] = p; sgv[k - 1] = sg; }
  }
  __syncthreads();
}

template <typename T, typename IdT, int NT>
__global__ void __launch_bounds__(EMB_THREADS)
embed_bwd_segment_kernel(const T* __restrict__ de, const IdT* __restrict__ ids, const IdT* __restrict__ seg,
                         const int* __restrict__ order, float* __restrict__ dW, float* __restrict__ partW,
                         float* __restrict__ partT, int M, int H, int vocab) {
  __shared__ long long sid[ESEG + 2];
  __shared__ int pos[ESEG], sgv[ESEG];
  const int s = blockIdx.x;
  load_segment_ids<IdT>(ids, NT > 0 ? seg : nullptr, order, s, M, vocab, sid, pos, sgv);
  const int vpr = H >> 3;
  for (int c = threadIdx.x; c < vpr; c += EMB_THREADS) {
    uint4 v[ESEG];
#pragma unroll
    for (int u = 0; u < ESEG; ++u) {
      const int p = pos[u];
      v[u] = p >= 0 ? reinterpret_cast<const uint4*>(de + static_cast<size_t>(p) * H)[c] : make_uint4(0, 0, 0, 0);
    }
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float tacc[NT > 0 ? NT : 1][8];
#pragma unroll
    for (int t = 0; t < (NT > 0 ? NT : 1); ++t)
#pragma unroll
      for (int k = 0; k < 8; ++k) tacc[t][k] = 0.f;
    int a = 0;                                  // first slot (within the segment) of the current run
#pragma unroll
    for (int u = 0; u < ESEG; ++u) {
      float f[8];
      unpack8e<T>(v[u], f);                     // T is 16-bit: zeros for a missing row
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
      if (NT > 0) {
        const int sg = sgv[u];
#pragma unroll
        for (int t = 0; t < NT; ++t)
          if (sg == t) {
#pragma unroll
            for (int k = 0; k < 8; ++k) tacc[t][k] += f[k];
          }
      }
      const long long id = sid[u + 1];
      if (u == ESEG - 1 || sid[u + 2] != id) {  // the run ends here, or leaves the segment
        if (id >= 0) {
          const bool enters = a == 0 && sid[0] == id;
          const bool leaves = u == ESEG - 1 && sid[ESEG + 1] == id;
          if (!enters && !leaves) {
            float4* dst = reinterpret_cast<float4*>(dW + static_cast<size_t>(id) * H + c * 8);
            float4 x = dst[0], y = dst[1];     // the decoder's weight gradient is already in there (tied matrix)
            x.x += acc[0]; x.y += acc[1]; x.z += acc[2]; x.w += acc[3];
            y.x += acc[4]; y.y += acc[5]; y.z += acc[6]; y.w += acc[7];
            dst[0] = x;
            dst[1] = y;
          } else {
            float4* dst = reinterpret_cast<float4*>(partW + (static_cast<size_t>(s) * 2 + (enters ? 0 : 1)) * H + c * 8);
            dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
            dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = 0.f;
        a = u + 1;
      }
    }
    if (NT > 0) {
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        float4* dst = reinterpret_cast<float4*>(partT + (static_cast<size_t>(s) * NT + t) * H + c * 8);
        dst[0] = make_float4(tacc[t][0], tacc[t][1], tacc[t][2], tacc[t][3]);
        dst[1] = make_float4(tacc[t][4], tacc[t][5], tacc[t][6], tacc[t][7]);
      }
    }
  }
}

template <typename IdT, int NT>
__global__ void __launch_bounds__(EMB_THREADS)
embed_bwd_combine_kernel(const IdT* __restrict__ ids, const int* __restrict__ order, const int* __restrict__ run_end,
                         const float* __restrict__ partW, const float* __restrict__ partT, float* __restrict__ dW,
                         float* __restrict__ dT, int accumulate_T, int M, int H, int vocab, int nseg) {
  if (static_cast<int>(blockIdx.x) >= nseg) {
    // ---- token-type sums: TYPE_COLS columns per CTA, segments dealt round-robin to TYPE_GROUPS thread groups
    if (NT == 0) return;
    __shared__ float red[TYPE_GROUPS][TYPE_COLS];
    const int col = (blockIdx.x - nseg) * TYPE_COLS + (threadIdx.x % TYPE_COLS), g = threadIdx.x / TYPE_COLS;
    for (int t = 0; t < NT; ++t) {
      float a = 0.f;
      if (col < H) {
#pragma unroll 8
        for (int s = g; s < nseg; s += TYPE_GROUPS) a += partT[(static_cast<size_t>(s) * NT + t) * H + col];
      }
      red[g][threadIdx.x % TYPE_COLS] = a;
      __syncthreads();
      if (g == 0 && col < H) {
        float r = red[0][threadIdx.x];
#pragma unroll
        for (int i = 1; i < TYPE_GROUPS; ++i) r += red[i][threadIdx.x];
        float* o = dT + static_cast<size_t>(t) * H + col;
        *o = accumulate_T ? *o + r : r;
      }
      __syncthreads();
    }
    return;
  }
  // ---- runs that cross segment boundaries: the segment where the run starts is its owner
  const int s = blockIdx.x;
  if ((s + 1) * ESEG >= M) return;             // nothing follows the last segment (uniform for the CTA)
  __shared__ long long sid[ESEG + 2];
  load_segment_ids<IdT>(ids, nullptr, order, s, M, vocab, sid, nullptr, nullptr);
  const long long id = sid[ESEG];              // the segment's last slot
  if (id < 0 || sid[ESEG + 1] != id) return;   // no run leaves this segment
  int a = ESEG - 1;
  while (a > 0 && sid[a] == id) --a;           // sid[a] is slot a - 1
  if (a == 0 && sid[0] == id) return;          // the run came in from the left: an earlier segment owns it
  const int e = min(run_end[s * ESEG + a], M);
  const int s_last = (e - 1) / ESEG;           // last segment the run reaches
  if (s_last <= s) return;                     // plan and ids disagree: leave dW alone rather than guess
  const int vpr = H >> 3;
  constexpr int U = 8;
  for (int c = threadIdx.x; c < vpr; c += EMB_THREADS) {
    const float4* first = reinterpret_cast<const float4*>(partW + (static_cast<size_t>(s) * 2 + 1) * H + c * 8);
    float4 x = first[0], y = first[1];
    for (int s2 = s + 1; s2 <= s_last; s2 += U) {
      float4 px[U], py[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (s2 + u <= s_last) {
          const float4* src = reinterpret_cast<const float4*>(partW + static_cast<size_t>(s2 + u) * 2 * H + c * 8);
          px[u] = src[0];
          py[u] = src[1];
        } else {
          px[u] = py[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        x.x += px[u].x; x.y += px[u].y; x.z += px[u].z; x.w += px[u].w;
        y.x += py[u].x; y.y += py[u].y; y.z += py[u].z; y.w += py[u].w;
      }
    }
    float4* dst = reinterpret_cast<float4*>(dW + static_cast<size_t>(id) * H + c * 8);
    float4 o0 = dst[0], o1 = dst[1];
    o0.x += x.x; o0.y += x.y; o0.z += x.z; o0.w += x.w;
    o1.x += y.x; o1.y += y.y; o1.z += y.z; o1.w += y.w;
    dst[0] = o0;
    dst[1] = o1;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
sum_over_batch_kernel(const T* __restrict__ x, int B, long long inner, float* __restrict__ out, int accumulate) {
  const long long j8 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (j8 * 8 >= inner) return;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 8
  for (int b = 0; b < B; ++b) {
    float f[8];
    unpack8e<T>(*reinterpret_cast<const uint4*>(x + static_cast<size_t>(b) * inner + j8 * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] += f[k];
  }
  float* o = out + j8 * 8;
#pragma unroll
  for (int k = 0; k < 8; ++k) o[k] = accumulate ? o[k] + acc[k] : acc[k];
}

template <typename T, typename IdT>
__global__ void __launch_bounds__(256)
colsum_where_partial_kernel(const T* __restrict__ x, const IdT* __restrict__ sel, long long want, int M, int N,
                            int rows_per_block, float* __restrict__ part) {
  const int c8 = blockIdx.x * blockDim.x + threadIdx.x;
  if (c8 * 8 >= N) return;
  const int r0 = blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int r = r0; r < r1; ++r) {
    if (static_cast<long long>(sel[r]) != want) continue;
    float f[8];
    unpack8e<T>(*reinterpret_cast<const uint4*>(x + static_cast<size_t>(r) * N + c8 * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] += f[k];
  }
  float* dst = part + static_cast<size_t>(blockIdx.y) * N + c8 * 8;
#pragma unroll
  for (int k = 0; k < 8; ++k) dst[k] = acc[k];
}

__global__ void __launch_bounds__(256)
colsum_parts2_kernel(const float* __restrict__ part, int nparts, int N, float* __restrict__ out, int accumulate) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= N) return;
  float a = 0.f;
  for (int i = 0; i < nparts; ++i) a += part[static_cast<size_t>(i) * N + h];
  out[h] = accumulate ? out[h] + a : a;
}

}  // namespace hs

using namespace hs;

extern "C" int hs_embed_gather_sum(const void* W, const void* P, const void* Tt, const void* ids, const void* seg,
                                   int ids_i64, void* z, int M, int S, int H, int vocab, int ntypes, int dtype,
                                   hs_stream_t stream) {
  if (M <= 0 || S <= 0 || H % 8 != 0) { set_last_error("hs_embed_gather_sum: bad shape M=%d S=%d H=%d", M, S, H); return 1; }
  const long long total = static_cast<long long>(M) * (H / 8);
  const int grid = static_cast<int>((total + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_EG(T)                                                                                                          \
  do {                                                                                                                    \
    if (ids_i64) embed_gather_sum_kernel<T, long long><<<grid, 256, 0, s>>>((const T*)W, (const T*)P, (const T*)Tt, (const long long*)ids, (const long long*)seg, (T*)z, M, S, H, vocab, ntypes); \
    else embed_gather_sum_kernel<T, int><<<grid, 256, 0, s>>>((const T*)W, (const T*)P, (const T*)Tt, (const int*)ids, (const int*)seg, (T*)z, M, S, H, vocab, ntypes); \
  } while (0)
  if (dtype == HS_F16) HS_EG(__half);
  else if (dtype == HS_BF16) HS_EG(__nv_bfloat16);
  else { set_last_error("hs_embed_gather_sum: dtype must be f16 or bf16"); return 1; }
#undef HS_EG
  return post_launch("embed_gather_sum_kernel");
}

extern "C" int hs_embed_scatter_word(const void* de, const void* ids, int ids_i64, float* dW, int M, int H, int vocab,
                                     int dtype, hs_stream_t stream) {
  if (M <= 0 || H % 8 != 0) { set_last_error("hs_embed_scatter_word: bad shape"); return 1; }
  const long long total = static_cast<long long>(M) * (H / 8);
  const int grid = static_cast<int>((total + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_ES(T)                                                                                                  \
  do {                                                                                                            \
    if (ids_i64) embed_scatter_word_kernel<T, long long><<<grid, 256, 0, s>>>((const T*)de, (const long long*)ids, dW, M, H, vocab); \
    else embed_scatter_word_kernel<T, int><<<grid, 256, 0, s>>>((const T*)de, (const int*)ids, dW, M, H, vocab);    \
  } while (0)
  if (dtype == HS_F16) HS_ES(__half);
  else if (dtype == HS_BF16) HS_ES(__nv_bfloat16);
  else { set_last_error("hs_embed_scatter_word: dtype must be f16 or bf16"); return 1; }
#undef HS_ES
  return post_launch("embed_scatter_word_kernel");
}

extern "C" long long hs_embed_backward_scratch_elems(int M, int H) {
  const long long nseg = (static_cast<long long>(M) + ESEG - 1) / ESEG;
  return nseg * 4 * H;                         // partW [nseg][2][H] + partT [nseg][2][H]
}

template <typename T, typename IdT, int NT>
static int launch_embed_backward(const void* de, const void* ids, const void* seg, const int* order, const int* run_end,
                                 float* dW, float* dT, int accumulate_T, float* scratch, int M, int H, int vocab,
                                 cudaStream_t s) {
  const int nseg = (M + ESEG - 1) / ESEG;
  float* partW = scratch;
  float* partT = scratch + static_cast<size_t>(nseg) * 2 * H;
  embed_bwd_segment_kernel<T, IdT, NT><<<nseg, EMB_THREADS, 0, s>>>(
      static_cast<const T*>(de), static_cast<const IdT*>(ids), static_cast<const IdT*>(seg), order, dW, partW, partT, M, H, vocab);
  if (post_launch("embed_bwd_segment_kernel")) return 1;
  const int type_ctas = NT > 0 ? (H + TYPE_COLS - 1) / TYPE_COLS : 0;
  embed_bwd_combine_kernel<IdT, NT><<<nseg + type_ctas, EMB_THREADS, 0, s>>>(
      static_cast<const IdT*>(ids), order, run_end, partW, partT, dW, dT, accumulate_T, M, H, vocab, nseg);
  return post_launch("embed_bwd_combine_kernel");
}

extern "C" int hs_embed_backward_sorted(const void* de, const void* ids, const void* seg, int ids_i64, const int* order,
                                        const int* run_end, float* dW, float* dT, int accumulate_T, float* scratch,
                                        int M, int H, int vocab, int dtype, hs_stream_t stream) {
  if (M <= 0 || H % 8 != 0 || order == nullptr || run_end == nullptr || scratch == nullptr || dW == nullptr) {
    set_last_error("hs_embed_backward_sorted: bad arguments");
    return 1;
  }
  if ((seg == nullptr) != (dT == nullptr)) { set_last_error("hs_embed_backward_sorted: seg and dT go together"); return 1; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool types = seg != nullptr;
#define HS_EB(T)                                                                                                         \
  do {                                                                                                                   \
    if (ids_i64) return types ? launch_embed_backward<T, long long, 2>(de, ids, seg, order, run_end, dW, dT, accumulate_T, scratch, M, H, vocab, s) \
                              : launch_embed_backward<T, long long, 0>(de, ids, seg, order, run_end, dW, dT, accumulate_T, scratch, M, H, vocab, s); \
    return types ? launch_embed_backward<T, int, 2>(de, ids, seg, order, run_end, dW, dT, accumulate_T, scratch, M, H, vocab, s) \
                 : launch_embed_backward<T, int, 0>(de, ids, seg, order, run_end, dW, dT, accumulate_T, scratch, M, H, vocab, s); \
  } while (0)
  if (dtype == HS_F16) HS_EB(__half);
  if (dtype == HS_BF16) HS_EB(__nv_bfloat16);
#undef HS_EB
  set_last_error("hs_embed_backward_sorted: dtype must be f16 or bf16");
  return 1;
}

extern "C" int hs_sum_over_batch(const void* x, int B, long long inner, float* out, int accumulate, int dtype,
                                 hs_stream_t stream) {
  if (B <= 0 || inner <= 0 || inner % 8 != 0) { set_last_error("hs_sum_over_batch: inner must be a multiple of 8"); return 1; }
  const int grid = static_cast<int>((inner / 8 + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (dtype == HS_F16) sum_over_batch_kernel<__half><<<grid, 256, 0, s>>>((const __half*)x, B, inner, out, accumulate);
  else if (dtype == HS_BF16) sum_over_batch_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, B, inner, out, accumulate);
  else { set_last_error("hs_sum_over_batch: dtype must be f16 or bf16"); return 1; }
  return post_launch("sum_over_batch_kernel");
}

extern "C" int hs_colsum_rows_where(const void* x, const void* sel, int sel_i64, long long want, int M, int N, float* out,
                                    int accumulate, float* scratch, int dtype, hs_stream_t stream) {
  if (M <= 0 || N <= 0 || N % 8 != 0 || scratch == nullptr) { set_last_error("hs_colsum_rows_where: bad arguments"); return 1; }
  const int slabs = colsum_slabs(M);
  const int rows_per_block = (M + slabs - 1) / slabs;
  dim3 grid((N / 8 + 255) / 256, slabs);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_CW(T)                                                                                                   \
  do {                                                                                                             \
    if (sel_i64) colsum_where_partial_kernel<T, long long><<<grid, 256, 0, s>>>((const T*)x, (const long long*)sel, want, M, N, rows_per_block, scratch); \
    else colsum_where_partial_kernel<T, int><<<grid, 256, 0, s>>>((const T*)x, (const int*)sel, want, M, N, rows_per_block, scratch); \
  } while (0)
  if (dtype == HS_F16) HS_CW(__half);
  else if (dtype == HS_BF16) HS_CW(__nv_bfloat16);
  else { set_last_error("hs_colsum_rows_where: dtype must be f16 or bf16"); return 1; }
#undef HS_CW
  if (post_launch("colsum_where_partial_kernel")) return 1;
  colsum_parts2_kernel<<<(N + 255) / 256, 256, 0, s>>>(scratch, slabs, N, out, accumulate);
  return post_launch("colsum_parts2_kernel");
}
