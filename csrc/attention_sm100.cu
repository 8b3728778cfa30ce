// hetseq_b200 -- fused masked-softmax attention on tcgen05 (forward and backward), head_dim 64.
//
// Reference: BertSelfAttention.forward (bert_modeling.py:351-377): scores = Q K^T / sqrt(d) + mask,
// softmax, dropout on the probabilities, P V, head merge -- six library launches that materialise the
// [B,h,S,S] score tensor three times. Here one CTA owns one (batch, head) pair:
//
//   forward   TMA: Q,K,V tiles [128 x 64] straight out of the packed [M,3H] projection (no head-split
//             copy) -> tcgen05.mma S = Q K^T into TMEM -> each of 128 threads owns one query row:
//             scale + additive mask + softmax + Philox dropout in registers -> P (16-bit) to swizzled
//             smem -> tcgen05.mma O = P V -> normalise, write [M,H] at the head's columns (head merge
//             for free) + log-sum-exp per row.
//   backward  S and dP = dO V^T by tcgen05 -> per row: P = exp(s - lse), the same dropout mask recomputed,
//             dS = P (dP - rowsum(dO O)) / sqrt(d) -> P and dS staged once in smem and consumed three
//             times through different UMMA descriptors (K-major for dQ = dS K, MN-major for
//             dV = P^T dO and dK = dS^T Q) -> dQ,dK,dV written into the packed [M,3H] gradient.
//
// The score matrix never leaves the SM. This file holds the single-tile kernels (sequence length 128, BERT
// phase 1); every other sequence length that is a multiple of 16 up to 512 is handled by
// attention_multi_sm100.cu (1..4 tiles, partial last tile), to which the host entry points below dispatch.
// Anything else (head_dim != 64, S > 512, S % 16 != 0) is reported as unsupported by hs_attention_supported().
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

#include <cstdlib>
#include <type_traits>

namespace hs {

constexpr int AT_S = 128;   // queries and keys per tile
constexpr int AT_D = 64;    // head dim
constexpr int AT_THREADS = 256;   // two threads per query row (column halves)
constexpr int TILE_BYTES = AT_S * 128;       // [128 rows x 64 x 16-bit], 128B rows, SW128
constexpr int PBUF_BYTES = 2 * TILE_BYTES;   // [128 x 128] 16-bit as two 64-column SW128 regions

struct AttnParams {
  const float* mask;      // [B, S] additive
  void* out;              // fwd: ctx [M, H];  bwd: dqkv [M, 3H]
  const void* o;          // bwd: forward output [M, H]
  const void* dout;       // bwd: dO [M, H] (also via tensor map)
  float* lse;             // [B*heads*S]
  float* dbias;           // bwd, optional: fp32 [3H] += column sums of dqkv (the q/k/v bias gradient)
  int B, heads, H;
  float scale, p;
  const long long* seed;
  uint32_t site;
};

__device__ __forceinline__ uint32_t sw128_offset(int row, int col16 /*16-byte chunk 0..7*/) {
  return static_cast<uint32_t>(row * 128 + ((col16 ^ (row & 7)) << 4));
}

template <typename T>
__device__ __forceinline__ void store_row_chunk8(uint8_t* buf, int row, int col0, const float (&v)[8]) {
  // 8 consecutive columns starting at col0 (multiple of 8) of a [128 x 128] two-region SW128 buffer
  const int region = col0 >> 6, c16 = (col0 & 63) >> 3;
  uint4 u;
  u.x = Cvt<T>::pack2(v[0], v[1]); u.y = Cvt<T>::pack2(v[2], v[3]);
  u.z = Cvt<T>::pack2(v[4], v[5]); u.w = Cvt<T>::pack2(v[6], v[7]);
  *reinterpret_cast<uint4*>(buf + region * TILE_BYTES + sw128_offset(row, c16)) = u;
}

// ------------------------------------------------------------------------------------------------
// 256 threads: warp w reads TMEM lane quarter (w & 3) -- query rows 32*(w&3) .. +31 -- and the column half
// (w >> 2); the two threads of a row meet through 1 KB of shared memory for the row max and the row sum.
template <typename T>
__global__ void __launch_bounds__(AT_THREADS, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + TILE_BYTES;
  uint8_t* sV = sK + TILE_BYTES;
  uint8_t* sP = sV + TILE_BYTES;
  float* sMask = reinterpret_cast<float*>(sP + PBUF_BYTES);
  float* sRed = sMask + AT_S;                       // [2][128] row max / row sum halves
  uint64_t* bars = reinterpret_cast<uint64_t*>(sRed + 2 * AT_S);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = warp & 3, hf = warp >> 2;
  const int r = q * 32 + lane;                      // query row of this thread
  const int b = blockIdx.x / P.heads, h = blockIdx.x % P.heads;
  const int row0 = b * AT_S;

  if (tid == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(&bars[2], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 256);
  }
  pdl_sync();   // set-up above overlapped the previous kernel's tail; global memory only from here on
  if (tid < AT_S) sMask[tid] = P.mask[b * AT_S + tid] * 1.4426950408889634f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (tid == 0) {
    mbar_expect_tx(&bars[0], 3 * TILE_BYTES);
    tma_load_2d(sQ, &tmQKV, &bars[0], h * AT_D, row0);
    tma_load_2d(sK, &tmQKV, &bars[0], P.H + h * AT_D, row0);
    tma_load_2d(sV, &tmQKV, &bars[0], 2 * P.H + h * AT_D, row0);
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    // S[q,k] = Q K^T : M=128, N=128, K=64
    const uint32_t idesc = umma_idesc_f16(sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value ? 1 : 0, 0, 0, 128, 128);
    const uint64_t aq = umma_desc_sw128(smem_u32(sQ), 16, 1024);
    const uint64_t bk = umma_desc_sw128(smem_u32(sK), 16, 1024);
#pragma unroll
    for (int k = 0; k < AT_D / 16; ++k) umma_f16(tmem, aq + k * 2, bk + k * 2, idesc, k != 0);
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 0);
  __syncwarp();
  tc_fence_after();

  // ---- softmax: this thread owns 64 key columns of query row r ----
  constexpr int HC = AT_S / 2;
  const int col0 = hf * HC;
  const uint32_t t_row = tmem + (static_cast<uint32_t>(q * 32) << 16);
  const float sl2 = P.scale * 1.4426950408889634f;   // fold log2(e) into the scale
  float s[HC];
#pragma unroll
  for (int c = 0; c < HC / 32; ++c) {
    uint32_t rr[32];
    tmem_ld_32x32(t_row + col0 + c * 32, rr);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(rr[i]) * sl2 + sMask[col0 + c * 32 + i];
  }
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < HC; ++i) mx = fmaxf(mx, s[i]);
  sRed[hf * AT_S + r] = mx;
  __syncthreads();
  mx = fmaxf(mx, sRed[(hf ^ 1) * AT_S + r]);
  __syncthreads();
  float l = 0.f;
#pragma unroll
  for (int i = 0; i < HC; ++i) { s[i] = exp2f(s[i] - mx); l += s[i]; }
  sRed[hf * AT_S + r] = l;
  const bool drop = P.p > 0.f;
  const float inv_keep = drop ? 1.f / (1.f - P.p) : 1.f;
  const uint64_t seed = drop ? philox_seed(P.seed) : 0ull;
  const uint32_t thr = dropout_thr16(P.p);
  const uint64_t e_row = (static_cast<uint64_t>(blockIdx.x) * AT_S + r) * AT_S;
#pragma unroll
  for (int c8 = 0; c8 < HC / 8; ++c8) {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = s[c8 * 8 + i];
    if (drop) {
      float mk[8];
      philox_keep8(seed, P.site, (e_row + col0 + c8 * 8) >> 3, thr, inv_keep, mk);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] *= mk[i];
    }
    store_row_chunk8<T>(sP, r, col0 + c8 * 8, v);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  l += sRed[(hf ^ 1) * AT_S + r];
  if (P.lse && hf == 0) P.lse[static_cast<size_t>(blockIdx.x) * AT_S + r] = (mx + log2f(l)) * 0.6931471805599453f;

  if (tid == 0) {
    tc_fence_after();
    // O[q,d] = P V : M=128, N=64, K=128; A = P (K-major, two 64-wide k regions), B = V (MN-major)
    const int fmt = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
    const uint32_t idesc = umma_idesc_f16(fmt, 0, 1, 128, 64);
#pragma unroll
    for (int kk = 0; kk < AT_S / 16; ++kk) {
      const uint64_t ap = umma_desc_sw128(smem_u32(sP) + (kk >> 2) * TILE_BYTES + (kk & 3) * 32, 16, 1024);
      const uint64_t bv = umma_desc_sw128(smem_u32(sV) + kk * 2048, TILE_BYTES, 1024);
      umma_f16(tmem + 128, ap, bv, idesc, kk != 0);
    }
    umma_commit(&bars[2]);
  }
  mbar_wait(&bars[2], 0);
  __syncwarp();
  tc_fence_after();
  {
    const float inv_l = 1.f / l;
    T* dst = reinterpret_cast<T*>(P.out) + static_cast<size_t>(row0 + r) * P.H + h * AT_D + hf * 32;
    uint32_t rr[32];
    tmem_ld_32x32(t_row + 128 + hf * 32, rr);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 u;
      u.x = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 0]) * inv_l, __uint_as_float(rr[i * 8 + 1]) * inv_l);
      u.y = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 2]) * inv_l, __uint_as_float(rr[i * 8 + 3]) * inv_l);
      u.z = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 4]) * inv_l, __uint_as_float(rr[i * 8 + 5]) * inv_l);
      u.w = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 6]) * inv_l, __uint_as_float(rr[i * 8 + 7]) * inv_l);
      reinterpret_cast<uint4*>(dst)[i] = u;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------------
// Backward, 256 threads, two CTAs per SM (96 KB of shared memory and 256 TMEM columns each):
//   phase 0  TMA Q,K,V,dO -> S = Q K^T (TMEM 0..127), dP = dO V^T (128..255)
//   phase 1  thread (row, column half): P = exp2(s - lse) with the dropout mask recomputed -> shared memory;
//            dS = P (dP - D) / sqrt(d) stays in registers as 32 packed 16-bit pairs
//   phase 2  dV = P^T dO into TMEM 0..63 (S is dead); when it has retired the same 32 KB buffer is
//            overwritten with dS and dK = dS^T Q (64..127), dQ = dS K (128..191) follow
//   phase 3  dQ,dK,dV to the packed [M,3H] gradient; their column sums -- the q/k/v bias gradient -- are
//            folded over the warp by a butterfly, over the CTA in shared memory, then 192 REDs per CTA.
template <typename T>
__global__ void __launch_bounds__(AT_THREADS, 2)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                const AttnParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + TILE_BYTES;
  uint8_t* sV = sK + TILE_BYTES;
  uint8_t* sDO = sV + TILE_BYTES;
  uint8_t* sPD = sDO + TILE_BYTES;            // P, later dS
  float* sMask = reinterpret_cast<float*>(sPD + PBUF_BYTES);
  float* sBias = sMask + AT_S;                // [3][64] column sums of dQ, dK, dV
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBias + 3 * AT_D);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = warp & 3, hf = warp >> 2;
  const int r = q * 32 + lane;
  const int b = blockIdx.x / P.heads, h = blockIdx.x % P.heads;
  const int row0 = b * AT_S;
  const int fmt = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;

  if (tid == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(&bars[2], 1);
    mbar_init(&bars[3], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 256);
  }
  pdl_sync();   // set-up above overlapped the previous kernel's tail; global memory only from here on
  if (tid < AT_S) sMask[tid] = P.mask[b * AT_S + tid] * 1.4426950408889634f;
  if (tid < 3 * AT_D) sBias[tid] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (tid == 0) {
    mbar_expect_tx(&bars[0], 4 * TILE_BYTES);
    tma_load_2d(sQ, &tmQKV, &bars[0], h * AT_D, row0);
    tma_load_2d(sK, &tmQKV, &bars[0], P.H + h * AT_D, row0);
    tma_load_2d(sV, &tmQKV, &bars[0], 2 * P.H + h * AT_D, row0);
    tma_load_2d(sDO, &tmDO, &bars[0], h * AT_D, row0);
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    const uint32_t idesc = umma_idesc_f16(fmt, 0, 0, 128, 128);
    const uint64_t aq = umma_desc_sw128(smem_u32(sQ), 16, 1024);
    const uint64_t bk = umma_desc_sw128(smem_u32(sK), 16, 1024);
    const uint64_t ado = umma_desc_sw128(smem_u32(sDO), 16, 1024);
    const uint64_t bv = umma_desc_sw128(smem_u32(sV), 16, 1024);
#pragma unroll
    for (int k = 0; k < AT_D / 16; ++k) umma_f16(tmem, aq + k * 2, bk + k * 2, idesc, k != 0);          // S = Q K^T
#pragma unroll
    for (int k = 0; k < AT_D / 16; ++k) umma_f16(tmem + 128, ado + k * 2, bv + k * 2, idesc, k != 0);   // dP = dO V^T
    umma_commit(&bars[1]);
  }

  // while the MMAs run: D = rowsum(dO * O) for this thread's query row, straight from global memory
  float Dq = 0.f;
  {
    const T* o = reinterpret_cast<const T*>(P.o) + static_cast<size_t>(row0 + r) * P.H + h * AT_D;
    const T* d = reinterpret_cast<const T*>(P.dout) + static_cast<size_t>(row0 + r) * P.H + h * AT_D;
#pragma unroll
    for (int i = 0; i < AT_D / 8; ++i) {
      const uint4 uo = reinterpret_cast<const uint4*>(o)[i], ud = reinterpret_cast<const uint4*>(d)[i];
      const float2 o0 = Cvt<T>::unpack2(uo.x), o1 = Cvt<T>::unpack2(uo.y), o2 = Cvt<T>::unpack2(uo.z), o3 = Cvt<T>::unpack2(uo.w);
      const float2 d0 = Cvt<T>::unpack2(ud.x), d1 = Cvt<T>::unpack2(ud.y), d2 = Cvt<T>::unpack2(ud.z), d3 = Cvt<T>::unpack2(ud.w);
      Dq += o0.x * d0.x + o0.y * d0.y + o1.x * d1.x + o1.y * d1.y + o2.x * d2.x + o2.y * d2.y + o3.x * d3.x + o3.y * d3.y;
    }
  }
  const float lse2 = P.lse[static_cast<size_t>(blockIdx.x) * AT_S + r] * 1.4426950408889634f;
  mbar_wait(&bars[1], 0);
  __syncwarp();
  tc_fence_after();

  constexpr int HC = AT_S / 2;
  const int col0 = hf * HC;
  const uint32_t t_row = tmem + (static_cast<uint32_t>(q * 32) << 16);
  const float sl2 = P.scale * 1.4426950408889634f;
  const bool drop = P.p > 0.f;
  const float inv_keep = drop ? 1.f / (1.f - P.p) : 1.f;
  const uint64_t seed = drop ? philox_seed(P.seed) : 0ull;
  const uint32_t thr = dropout_thr16(P.p);
  const uint64_t e_row = (static_cast<uint64_t>(blockIdx.x) * AT_S + r) * AT_S;
  uint32_t dsp[HC / 2];                        // dS of this thread's 64 columns, packed 16-bit pairs
#pragma unroll
  for (int c = 0; c < HC / 32; ++c) {
    uint32_t rs[32], rd[32];
    tmem_ld_32x32(t_row + col0 + c * 32, rs);
    tmem_ld_32x32(t_row + 128 + col0 + c * 32, rd);
    tmem_ld_wait();
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float pv[8], mk[8];
      if (drop) {
        philox_keep8(seed, P.site, (e_row + col0 + c * 32 + g * 8) >> 3, thr, inv_keep, mk);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) mk[i] = 1.f;
      }
      float dsv[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int col = col0 + c * 32 + g * 8 + i;
        const float p = exp2f(__uint_as_float(rs[g * 8 + i]) * sl2 + sMask[col] - lse2);
        const float dp = __uint_as_float(rd[g * 8 + i]) * mk[i];
        pv[i] = p * mk[i];                        // dropped probabilities (what multiplied V in forward)
        dsv[i] = p * (dp - Dq) * P.scale;         // d(scores) including the 1/sqrt(d)
      }
      store_row_chunk8<T>(sPD, r, col0 + c * 32 + g * 8, pv);
#pragma unroll
      for (int i = 0; i < 4; ++i) dsp[c * 16 + g * 4 + i] = Cvt<T>::pack2(dsv[2 * i], dsv[2 * i + 1]);
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();

  const uint32_t id_mm = umma_idesc_f16(fmt, 1, 1, 128, 64);   // A MN-major, B MN-major
  const uint32_t id_km = umma_idesc_f16(fmt, 0, 1, 128, 64);   // A K-major,  B MN-major
  if (tid == 0) {
    tc_fence_after();
#pragma unroll
    for (int kk = 0; kk < AT_S / 16; ++kk) {
      // reduction over queries: 16 query rows per step = 2048 B in every operand
      const uint64_t aP = umma_desc_sw128(smem_u32(sPD) + kk * 2048, TILE_BYTES, 1024);
      const uint64_t bDO = umma_desc_sw128(smem_u32(sDO) + kk * 2048, TILE_BYTES, 1024);
      umma_f16(tmem, aP, bDO, id_mm, kk != 0);          // dV[k,d] += P_drop^T dO
    }
    umma_commit(&bars[2]);
  }
  mbar_wait(&bars[2], 0);       // P has been consumed: the buffer can take dS
  __syncwarp();
  tc_fence_after();
#pragma unroll
  for (int c8 = 0; c8 < HC / 8; ++c8) {
    const int col = col0 + c8 * 8;
    const int region = col >> 6, c16 = (col & 63) >> 3;
    *reinterpret_cast<uint4*>(sPD + region * TILE_BYTES + sw128_offset(r, c16)) =
        make_uint4(dsp[c8 * 4 + 0], dsp[c8 * 4 + 1], dsp[c8 * 4 + 2], dsp[c8 * 4 + 3]);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
#pragma unroll
    for (int kk = 0; kk < AT_S / 16; ++kk) {
      const uint64_t aDS = umma_desc_sw128(smem_u32(sPD) + kk * 2048, TILE_BYTES, 1024);
      const uint64_t bQ = umma_desc_sw128(smem_u32(sQ) + kk * 2048, TILE_BYTES, 1024);
      umma_f16(tmem + 64, aDS, bQ, id_mm, kk != 0);     // dK[k,d] += dS^T Q
      // reduction over keys: dS K-major (two 64-wide regions), K tile MN-major
      const uint64_t aDSk = umma_desc_sw128(smem_u32(sPD) + (kk >> 2) * TILE_BYTES + (kk & 3) * 32, 16, 1024);
      const uint64_t bK = umma_desc_sw128(smem_u32(sK) + kk * 2048, TILE_BYTES, 1024);
      umma_f16(tmem + 128, aDSk, bK, id_km, kk != 0);   // dQ[q,d] += dS K
    }
    umma_commit(&bars[3]);
  }
  mbar_wait(&bars[3], 0);
  __syncwarp();
  tc_fence_after();
  {
    T* base = reinterpret_cast<T*>(P.out) + static_cast<size_t>(row0 + r) * (3 * P.H) + h * AT_D + hf * 32;
#pragma unroll
    for (int which = 0; which < 3; ++which) {
      // which: 0 -> dQ (TMEM 128), 1 -> dK (64), 2 -> dV (0); this thread: 32 of the 64 columns
      const uint32_t col = which == 0 ? 128u : (which == 1 ? 64u : 0u);
      T* dst = base + which * P.H;
      uint32_t rr[32];
      tmem_ld_32x32(t_row + col + hf * 32, rr);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 u;
        u.x = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 0]), __uint_as_float(rr[i * 8 + 1]));
        u.y = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 2]), __uint_as_float(rr[i * 8 + 3]));
        u.z = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 4]), __uint_as_float(rr[i * 8 + 5]));
        u.w = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 6]), __uint_as_float(rr[i * 8 + 7]));
        reinterpret_cast<uint4*>(dst)[i] = u;
      }
      if (P.dbias != nullptr) {
        float cs[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) cs[i] = __uint_as_float(rr[i]);
        warp_colsum<32>(cs, lane);
        atomicAdd(&sBias[which * AT_D + hf * 32 + lane], cs[0]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (P.dbias != nullptr && tid < 3 * AT_D) {
    const int which = tid / AT_D, c = tid % AT_D;
    atomicAdd(P.dbias + which * P.H + h * AT_D + c, sBias[tid]);
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

constexpr int ATTN_FWD_SMEM = 3 * TILE_BYTES + PBUF_BYTES + 3 * AT_S * 4 + 64 + 1024;
constexpr int ATTN_BWD_SMEM = 4 * TILE_BYTES + PBUF_BYTES + AT_S * 4 + 3 * AT_D * 4 + 64 + 1024;

// multi-tile kernels (attention_multi_sm100.cu): S = 256, 384, 512
bool attn_multi_supported(int S, int head_dim);
int attn_multi_fwd(const void* qkv, const float* mask, void* ctx, float* lse, int B, int S, int heads, int H,
                   float dropout_p, const long long* seed, unsigned site, int dtype, cudaStream_t s);
int attn_multi_bwd(const void* qkv, const float* mask, const void* ctx, const void* dctx, const float* lse,
                   void* dqkv, float* dbias, int B, int S, int heads, int H, float dropout_p, const long long* seed,
                   unsigned site, int dtype, cudaStream_t s);
// HETSEQ_ATTN_MULTI=0 keeps sequence lengths other than 128 on the caller's composite path
static bool multi_enabled() {
  static const bool on = [] { const char* e = getenv("HETSEQ_ATTN_MULTI"); return e == nullptr || atoi(e) != 0; }();
  return on;
}

// shared-memory opt-in and carve-out are function attributes PER DEVICE: do it once for every device used
template <typename K>
static int prepare_attn(K kern, int smem, bool (&done)[64]) {
  int dev = 0;
  HS_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) dev = 63;
  if (!done[dev]) {
    HS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    HS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    done[dev] = true;
  }
  return 0;
}

static int attn_check(const char* who, int B, int S, int heads, int H) {
  if (S != AT_S || heads <= 0 || H != heads * AT_D || B <= 0) {
    set_last_error("%s: unsupported shape B=%d S=%d heads=%d H=%d (needs S=%d, head_dim=%d)", who, B, S, heads, H, AT_S, AT_D);
    return 1;
  }
  return 0;
}

}  // namespace hs

using namespace hs;

extern "C" int hs_attention_supported(int S, int head_dim) {
  if (S == AT_S && head_dim == AT_D) return 1;
  return (multi_enabled() && attn_multi_supported(S, head_dim)) ? 1 : 0;
}

extern "C" int hs_attention_fwd(const void* qkv, const float* mask, void* ctx, float* lse, int B, int S, int heads,
                                int H, float dropout_p, const long long* seed, unsigned site, int dtype,
                                hs_stream_t stream) {
  if (dropout_p > 0.f && seed == nullptr) { set_last_error("hs_attention_fwd: dropout needs a seed pointer"); return 1; }
  if (S != AT_S && heads > 0 && B > 0 && H == heads * AT_D && attn_multi_supported(S, AT_D))
    return attn_multi_fwd(qkv, mask, ctx, lse, B, S, heads, H, dropout_p, seed, site, dtype, static_cast<cudaStream_t>(stream));
  if (attn_check("hs_attention_fwd", B, S, heads, H)) return 1;
  CUtensorMap tm;
  if (make_tmap_2d_16b(&tm, qkv, 3ull * H, static_cast<uint64_t>(B) * S, 3ull * H, 64, AT_S)) return 1;
  const float scale = 1.0f / sqrtf(static_cast<float>(H / heads));      // 1/sqrt(d), bert_modeling.py:361-362
  AttnParams P{mask, ctx, nullptr, nullptr, lse, nullptr, B, heads, H, scale, dropout_p, seed, site};
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = B * heads;
  if (dtype == HS_F16) {
    static bool done[64] = {false};
    if (prepare_attn(attn_fwd_kernel<__half>, ATTN_FWD_SMEM, done)) return 1;
    (void)launch_k(attn_fwd_kernel<__half>, dim3(grid), dim3(AT_THREADS), ATTN_FWD_SMEM, s, tm, P);
  } else if (dtype == HS_BF16) {
    static bool done[64] = {false};
    if (prepare_attn(attn_fwd_kernel<__nv_bfloat16>, ATTN_FWD_SMEM, done)) return 1;
    (void)launch_k(attn_fwd_kernel<__nv_bfloat16>, dim3(grid), dim3(AT_THREADS), ATTN_FWD_SMEM, s, tm, P);
  } else { set_last_error("hs_attention_fwd: dtype must be f16 or bf16"); return 1; }
  return post_launch("attn_fwd_kernel");
}

extern "C" int hs_attention_bwd(const void* qkv, const float* mask, const void* ctx, const void* dctx, const float* lse,
                                void* dqkv, float* dbias, int accumulate_bias, int B, int S, int heads, int H,
                                float dropout_p, const long long* seed, unsigned site, int dtype,
                                hs_stream_t stream) {
  const bool multi = S != AT_S && heads > 0 && B > 0 && H == heads * AT_D && attn_multi_supported(S, AT_D);
  if (!multi && attn_check("hs_attention_bwd", B, S, heads, H)) return 1;
  if (dbias != nullptr && !accumulate_bias)
    HS_CUDA_OK(cudaMemsetAsync(dbias, 0, sizeof(float) * 3 * H, static_cast<cudaStream_t>(stream)));
  if (multi) {
    if (dropout_p > 0.f && seed == nullptr) { set_last_error("hs_attention_bwd: dropout needs a seed pointer"); return 1; }
    return attn_multi_bwd(qkv, mask, ctx, dctx, lse, dqkv, dbias, B, S, heads, H, dropout_p, seed, site, dtype,
                          static_cast<cudaStream_t>(stream));
  }
  if (dropout_p > 0.f && seed == nullptr) { set_last_error("hs_attention_bwd: dropout needs a seed pointer"); return 1; }
  CUtensorMap tmQ, tmD;
  if (make_tmap_2d_16b(&tmQ, qkv, 3ull * H, static_cast<uint64_t>(B) * S, 3ull * H, 64, AT_S)) return 1;
  if (make_tmap_2d_16b(&tmD, dctx, static_cast<uint64_t>(H), static_cast<uint64_t>(B) * S, static_cast<uint64_t>(H), 64, AT_S)) return 1;
  const float scale = 1.0f / sqrtf(static_cast<float>(H / heads));
  AttnParams P{mask, dqkv, ctx, dctx, const_cast<float*>(lse), dbias, B, heads, H, scale, dropout_p, seed, site};
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = B * heads;
  if (dtype == HS_F16) {
    static bool done[64] = {false};
    if (prepare_attn(attn_bwd_kernel<__half>, ATTN_BWD_SMEM, done)) return 1;
    (void)launch_k(attn_bwd_kernel<__half>, dim3(grid), dim3(AT_THREADS), ATTN_BWD_SMEM, s, tmQ, tmD, P);
  } else if (dtype == HS_BF16) {
    static bool done[64] = {false};
    if (prepare_attn(attn_bwd_kernel<__nv_bfloat16>, ATTN_BWD_SMEM, done)) return 1;
    (void)launch_k(attn_bwd_kernel<__nv_bfloat16>, dim3(grid), dim3(AT_THREADS), ATTN_BWD_SMEM, s, tmQ, tmD, P);
  } else { set_last_error("hs_attention_bwd: dtype must be f16 or bf16"); return 1; }
  return post_launch("attn_bwd_kernel");
}
