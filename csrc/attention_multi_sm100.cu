// hetseq_b200 -- fused masked-softmax attention for any sequence length that is a multiple of 16 up to 512
// (1..4 tiles of 128 rows, the last one possibly partial; BERT phase 2: S = 512), head_dim 64, on tcgen05. Same math and the same Philox dropout stream as attention_sm100.cu
// (reference: bert_modeling.py:351-377), which keeps the single-tile S = 128 case.
//
//   forward    one CTA per (batch, head, query tile). Pass 1 walks the key tiles computing only the row maximum
//              (S_j = Q K_j^T in TMEM, no exp); pass 2 walks them again: P_j = exp2(s - max) with dropout ->
//              swizzled smem -> O += P_j V_j accumulated in TMEM. Knowing the true maximum up front means the
//              accumulator never has to be rescaled (no TMEM read-modify-write per tile); the price is a second
//              QK^T, which is a quarter of the tile's tensor work and far from the bottleneck.
//   backward   two launches of one kernel template, no atomics on the outputs:
//              MODE 0: one CTA per (batch, head, KEY tile j), loops over query tiles i, accumulates
//                      dV_j += P_ij^T dO_i and dK_j += dS_ij^T Q_i in TMEM;
//              MODE 1: one CTA per (batch, head, QUERY tile i), loops over key tiles j, accumulates
//                      dQ_i += dS_ij K_j in TMEM.
//              Both recompute S, dP = dO V^T, P and dS per (i, j) pair exactly like the single-tile kernel.
//
// The [S,S] score matrix never exists in global memory; HBM traffic is the packed QKV projection, O, dO and
// the gradients, each tile read n = ceil(S/128) times. Tiles come through 3-D tensor maps [B][S][cols]: the rows
// of a partial last tile beyond S are out of bounds and arrive as zeros (never the next sequence's rows); the
// padded key columns get -inf in the mask, the padded query rows p = 0 in backward and no stores anywhere.
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

#include <type_traits>

namespace hs {

namespace {

constexpr int MT = 128;                  // tile: queries and keys per step
constexpr int MD = 64;                   // head dim
constexpr int MTHREADS = 256;            // two threads per query row (column halves)
constexpr int MTILE_BYTES = MT * 128;    // [128 rows x 64 x 16-bit], 128 B rows, SW128
constexpr int MPBUF_BYTES = 2 * MTILE_BYTES;
constexpr int MAX_S = 512;
constexpr float LOG2E = 1.4426950408889634f;

struct MultiParams {
  const float* mask;      // [B, S] additive
  void* out;              // fwd: ctx [M, H];  bwd: dqkv [M, 3H]
  const void* o;          // bwd: forward output [M, H]
  const void* dout;       // bwd: dO [M, H]
  float* lse;             // [B*heads*S]
  float* dbias;           // bwd, optional: fp32 [3H]
  int B, heads, H, S, ntiles;
  float scale, p;
  const long long* seed;
  uint32_t site;
};

__device__ __forceinline__ uint32_t sw_off(int row, int col16) {
  return static_cast<uint32_t>(row * 128 + ((col16 ^ (row & 7)) << 4));
}

// 8 consecutive columns starting at col0 (multiple of 8) of a [128 x 128] two-region SW128 buffer
template <typename T>
__device__ __forceinline__ void put8(uint8_t* buf, int row, int col0, const float (&v)[8]) {
  const int region = col0 >> 6, c16 = (col0 & 63) >> 3;
  uint4 u;
  u.x = Cvt<T>::pack2(v[0], v[1]); u.y = Cvt<T>::pack2(v[2], v[3]);
  u.z = Cvt<T>::pack2(v[4], v[5]); u.w = Cvt<T>::pack2(v[6], v[7]);
  *reinterpret_cast<uint4*>(buf + region * MTILE_BYTES + sw_off(row, c16)) = u;
}

template <typename T> constexpr int fmt_of() { return std::is_same<T, __nv_bfloat16>::value ? 1 : 0; }

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(MTHREADS, 2)
attn_fwd_multi_kernel(const __grid_constant__ CUtensorMap tmQKV, const MultiParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + MTILE_BYTES;
  uint8_t* sV = sK + MTILE_BYTES;
  uint8_t* sP = sV + MTILE_BYTES;
  float* sMask = reinterpret_cast<float*>(sP + MPBUF_BYTES);     // [S] additive mask * log2(e)
  float* sRed = sMask + MAX_S;                                   // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sRed + 2 * MT);   // 0: TMA, 1: S ready, 2: O step done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = warp & 3, hf = warp >> 2;
  const int r = q * 32 + lane;
  const int n = P.ntiles;
  const int qi = blockIdx.x % n;
  const int bh = blockIdx.x / n;
  const int b = bh / P.heads, h = bh % P.heads;
  const int qrow0 = b * P.S + qi * MT;           // first row of this query tile in [M, *]

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
  for (int i = tid; i < n * MT; i += MTHREADS) sMask[i] = i < P.S ? P.mask[b * P.S + i] * LOG2E : -INFINITY;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t t_row = tmem + (static_cast<uint32_t>(q * 32) << 16);
  const uint32_t idesc_s = umma_idesc_f16(fmt_of<T>(), 0, 0, 128, 128);
  const uint32_t idesc_o = umma_idesc_f16(fmt_of<T>(), 0, 1, 128, 64);
  const float sl2 = P.scale * LOG2E;
  constexpr int HC = MT / 2;
  const int col0 = hf * HC;
  uint32_t ph_load = 0, ph_s = 0, ph_o = 0;

  // ---- pass 1: row maximum over all key tiles ----
  float mx = -INFINITY;
  for (int j = 0; j < n; ++j) {
    if (tid == 0) {
      mbar_expect_tx(&bars[0], (j == 0 ? 2 : 1) * MTILE_BYTES);
      if (j == 0) tma_load_3d(sQ, &tmQKV, &bars[0], h * MD, qi * MT, b);
      tma_load_3d(sK, &tmQKV, &bars[0], P.H + h * MD, j * MT, b);
      mbar_wait(&bars[0], ph_load);
      tc_fence_after();
      const uint64_t aq = umma_desc_sw128(smem_u32(sQ), 16, 1024);
      const uint64_t bk = umma_desc_sw128(smem_u32(sK), 16, 1024);
#pragma unroll
      for (int k = 0; k < MD / 16; ++k) umma_f16(tmem, aq + k * 2, bk + k * 2, idesc_s, k != 0);
      umma_commit(&bars[1]);
    }
    ph_load ^= 1;
    mbar_wait(&bars[1], ph_s);
    ph_s ^= 1;
    __syncwarp();
    tc_fence_after();
#pragma unroll
    for (int c = 0; c < HC / 32; ++c) {
      uint32_t rr[32];
      tmem_ld_32x32(t_row + col0 + c * 32, rr);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(rr[i]) * sl2 + sMask[j * MT + col0 + c * 32 + i]);
    }
    // every thread has read S (TMEM) and the MMA has consumed sK: both may be overwritten by the next step
    tc_fence_before();
    __syncthreads();
  }
  sRed[hf * MT + r] = mx;
  __syncthreads();
  mx = fmaxf(mx, sRed[(hf ^ 1) * MT + r]);
  __syncthreads();

  // ---- pass 2: probabilities, dropout, O += P V ----
  const bool drop = P.p > 0.f;
  const float inv_keep = drop ? 1.f / (1.f - P.p) : 1.f;
  const uint64_t seed = drop ? philox_seed(P.seed) : 0ull;
  const uint32_t thr = dropout_thr16(P.p);
  const uint64_t e_row = (static_cast<uint64_t>(bh) * P.S + qi * MT + r) * P.S;
  float l = 0.f;
  for (int j = 0; j < n; ++j) {
    if (tid == 0) {
      tc_fence_after();
      mbar_expect_tx(&bars[0], 2 * MTILE_BYTES);
      tma_load_3d(sK, &tmQKV, &bars[0], P.H + h * MD, j * MT, b);
      tma_load_3d(sV, &tmQKV, &bars[0], 2 * P.H + h * MD, j * MT, b);
      mbar_wait(&bars[0], ph_load);
      tc_fence_after();
      const uint64_t aq = umma_desc_sw128(smem_u32(sQ), 16, 1024);
      const uint64_t bk = umma_desc_sw128(smem_u32(sK), 16, 1024);
#pragma unroll
      for (int k = 0; k < MD / 16; ++k) umma_f16(tmem, aq + k * 2, bk + k * 2, idesc_s, k != 0);
      umma_commit(&bars[1]);
    }
    ph_load ^= 1;
    mbar_wait(&bars[1], ph_s);
    ph_s ^= 1;
    __syncwarp();
    tc_fence_after();
#pragma unroll
    for (int c = 0; c < HC / 32; ++c) {
      uint32_t rr[32];
      tmem_ld_32x32(t_row + col0 + c * 32, rr);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int col = col0 + c * 32 + g * 8 + i;
          v[i] = exp2f(__uint_as_float(rr[g * 8 + i]) * sl2 + sMask[j * MT + col] - mx);
          l += v[i];
        }
        if (drop) {
          float mk[8];
          philox_keep8(seed, P.site, (e_row + j * MT + col0 + c * 32 + g * 8) >> 3, thr, inv_keep, mk);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] *= mk[i];
        }
        put8<T>(sP, r, col0 + c * 32 + g * 8, v);
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < MT / 16; ++kk) {
        const uint64_t ap = umma_desc_sw128(smem_u32(sP) + (kk >> 2) * MTILE_BYTES + (kk & 3) * 32, 16, 1024);
        const uint64_t bv = umma_desc_sw128(smem_u32(sV) + kk * 2048, MTILE_BYTES, 1024);
        umma_f16(tmem + 128, ap, bv, idesc_o, (j != 0 || kk != 0) ? 1u : 0u);
      }
      umma_commit(&bars[2]);
    }
    mbar_wait(&bars[2], ph_o);       // P, V and K of this step are consumed
    ph_o ^= 1;
    __syncwarp();
    tc_fence_after();
  }

  sRed[hf * MT + r] = l;
  __syncthreads();
  l += sRed[(hf ^ 1) * MT + r];
  const bool row_ok = qi * MT + r < P.S;          // rows of a partial last tile beyond S own nothing
  if (P.lse && hf == 0 && row_ok)
    P.lse[static_cast<size_t>(bh) * P.S + qi * MT + r] = (mx + log2f(l)) * 0.6931471805599453f;
  {
    const float inv_l = 1.f / l;
    T* dst = reinterpret_cast<T*>(P.out) + static_cast<size_t>(qrow0 + r) * P.H + h * MD + hf * 32;
    uint32_t rr[32];
    tmem_ld_32x32(t_row + 128 + hf * 32, rr);      // warp-collective (.sync.aligned): never under a per-row condition
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 4 && row_ok; ++i) {
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
// backward. MODE 0: CTA = key tile (outer), loop over query tiles -> dK, dV.  MODE 1: CTA = query tile, loop
// over key tiles -> dQ. TMEM columns: S [0,128)  dP [128,256)  acc0 [256,320) (dV | dQ)  acc1 [320,384) (dK).
// ------------------------------------------------------------------------------------------------
template <typename T, int MODE>
__global__ void __launch_bounds__(MTHREADS, 1)
attn_bwd_multi_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                      const MultiParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + MTILE_BYTES;
  uint8_t* sV = sK + MTILE_BYTES;
  uint8_t* sDO = sV + MTILE_BYTES;
  uint8_t* sPD = sDO + MTILE_BYTES;            // P, then dS (MODE 0); dS (MODE 1)
  float* sMask = reinterpret_cast<float*>(sPD + MPBUF_BYTES);    // [S] additive mask * log2(e)
  float* sBias = sMask + MAX_S;                                  // [2][64] column sums of the two accumulators
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBias + 2 * MD);  // 0: TMA, 1: S/dP ready, 2: first MMA group, 3: second
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = warp & 3, hf = warp >> 2;
  const int r = q * 32 + lane;
  const int n = P.ntiles;
  const int outer = blockIdx.x % n;
  const int bh = blockIdx.x / n;
  const int b = bh / P.heads, h = bh % P.heads;
  constexpr int fmt = fmt_of<T>();

  if (tid == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
#pragma unroll
    for (int i = 0; i < 4; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  pdl_sync();   // set-up above overlapped the previous kernel's tail; global memory only from here on
  for (int i = tid; i < n * MT; i += MTHREADS) sMask[i] = i < P.S ? P.mask[b * P.S + i] * LOG2E : -INFINITY;
  if (tid < 2 * MD) sBias[tid] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t t_row = tmem + (static_cast<uint32_t>(q * 32) << 16);
  const uint32_t id_ss = umma_idesc_f16(fmt, 0, 0, 128, 128);
  const uint32_t id_mm = umma_idesc_f16(fmt, 1, 1, 128, 64);   // A MN-major, B MN-major
  const uint32_t id_km = umma_idesc_f16(fmt, 0, 1, 128, 64);   // A K-major,  B MN-major
  const float sl2 = P.scale * LOG2E;
  const bool drop = P.p > 0.f;
  const float inv_keep = drop ? 1.f / (1.f - P.p) : 1.f;
  const uint64_t seed = drop ? philox_seed(P.seed) : 0ull;
  const uint32_t thr = dropout_thr16(P.p);
  constexpr int HC = MT / 2;
  const int col0 = hf * HC;
  uint32_t ph_load = 0, ph_sdp = 0, ph_a = 0, ph_b = 0;

  // per-row quantities of a query tile: log-sum-exp and D = rowsum(dO * O)
  float lse2 = 0.f, Dq = 0.f;
  auto row_stats = [&](int qt) {
    if (qt * MT + r >= P.S) {      // padded query row: p = exp2(.. - inf) = 0, so it adds nothing to dK / dV
      Dq = 0.f;
      lse2 = INFINITY;
      return;
    }
    const size_t grow = static_cast<size_t>(b) * P.S + qt * MT + r;
    const T* o = reinterpret_cast<const T*>(P.o) + grow * P.H + h * MD;
    const T* d = reinterpret_cast<const T*>(P.dout) + grow * P.H + h * MD;
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < MD / 8; ++i) {
      const uint4 uo = reinterpret_cast<const uint4*>(o)[i], ud = reinterpret_cast<const uint4*>(d)[i];
      const float2 o0 = Cvt<T>::unpack2(uo.x), o1 = Cvt<T>::unpack2(uo.y), o2 = Cvt<T>::unpack2(uo.z), o3 = Cvt<T>::unpack2(uo.w);
      const float2 d0 = Cvt<T>::unpack2(ud.x), d1 = Cvt<T>::unpack2(ud.y), d2 = Cvt<T>::unpack2(ud.z), d3 = Cvt<T>::unpack2(ud.w);
      acc += o0.x * d0.x + o0.y * d0.y + o1.x * d1.x + o1.y * d1.y + o2.x * d2.x + o2.y * d2.y + o3.x * d3.x + o3.y * d3.y;
    }
    Dq = acc;
    lse2 = P.lse[static_cast<size_t>(bh) * P.S + qt * MT + r] * LOG2E;
  };
  if (MODE == 1) row_stats(outer);

  for (int it = 0; it < n; ++it) {
    const int qt = MODE == 0 ? it : outer;       // query tile of this step
    const int kt = MODE == 0 ? outer : it;       // key tile of this step
    if (tid == 0) {
      tc_fence_after();
      const int qrow = qt * MT, krow = kt * MT;      // row inside the sequence; the batch is the third coordinate
      const bool load_q = MODE == 0 || it == 0, load_k = MODE == 1 || it == 0;
      mbar_expect_tx(&bars[0], ((load_q ? 2 : 0) + (load_k ? 2 : 0)) * MTILE_BYTES);
      if (load_q) {
        tma_load_3d(sQ, &tmQKV, &bars[0], h * MD, qrow, b);
        tma_load_3d(sDO, &tmDO, &bars[0], h * MD, qrow, b);
      }
      if (load_k) {
        tma_load_3d(sK, &tmQKV, &bars[0], P.H + h * MD, krow, b);
        tma_load_3d(sV, &tmQKV, &bars[0], 2 * P.H + h * MD, krow, b);
      }
      mbar_wait(&bars[0], ph_load);
      tc_fence_after();
      const uint64_t aq = umma_desc_sw128(smem_u32(sQ), 16, 1024);
      const uint64_t bk = umma_desc_sw128(smem_u32(sK), 16, 1024);
      const uint64_t ado = umma_desc_sw128(smem_u32(sDO), 16, 1024);
      const uint64_t bv = umma_desc_sw128(smem_u32(sV), 16, 1024);
#pragma unroll
      for (int k = 0; k < MD / 16; ++k) umma_f16(tmem, aq + k * 2, bk + k * 2, id_ss, k != 0);          // S = Q K^T
#pragma unroll
      for (int k = 0; k < MD / 16; ++k) umma_f16(tmem + 128, ado + k * 2, bv + k * 2, id_ss, k != 0);   // dP = dO V^T
      umma_commit(&bars[1]);
    }
    ph_load ^= 1;
    if (MODE == 0) row_stats(qt);                // overlaps with the TMA + MMAs
    mbar_wait(&bars[1], ph_sdp);
    ph_sdp ^= 1;
    __syncwarp();
    tc_fence_after();

    const uint64_t e_row = (static_cast<uint64_t>(bh) * P.S + qt * MT + r) * P.S + static_cast<uint64_t>(kt) * MT;
    uint32_t dsp[MODE == 0 ? HC / 2 : 1];        // MODE 0: dS of this thread's 64 columns, packed 16-bit pairs
#pragma unroll
    for (int c = 0; c < HC / 32; ++c) {
      uint32_t rs[32], rd[32];
      tmem_ld_32x32(t_row + col0 + c * 32, rs);
      tmem_ld_32x32(t_row + 128 + col0 + c * 32, rd);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float mk[8];
        if (drop) {
          philox_keep8(seed, P.site, (e_row + col0 + c * 32 + g * 8) >> 3, thr, inv_keep, mk);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) mk[i] = 1.f;
        }
        float pv[8], dsv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int col = col0 + c * 32 + g * 8 + i;
          const float p = exp2f(__uint_as_float(rs[g * 8 + i]) * sl2 + sMask[kt * MT + col] - lse2);
          const float dp = __uint_as_float(rd[g * 8 + i]) * mk[i];
          pv[i] = p * mk[i];
          dsv[i] = p * (dp - Dq) * P.scale;
        }
        if (MODE == 0) {
          put8<T>(sPD, r, col0 + c * 32 + g * 8, pv);
#pragma unroll
          for (int i = 0; i < 4; ++i) dsp[MODE == 0 ? c * 16 + g * 4 + i : 0] = Cvt<T>::pack2(dsv[2 * i], dsv[2 * i + 1]);
        } else {
          put8<T>(sPD, r, col0 + c * 32 + g * 8, dsv);
        }
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();

    if (MODE == 0) {
      if (tid == 0) {
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < MT / 16; ++kk) {
          const uint64_t aP = umma_desc_sw128(smem_u32(sPD) + kk * 2048, MTILE_BYTES, 1024);
          const uint64_t bDO = umma_desc_sw128(smem_u32(sDO) + kk * 2048, MTILE_BYTES, 1024);
          umma_f16(tmem + 256, aP, bDO, id_mm, (it != 0 || kk != 0) ? 1u : 0u);      // dV[k,d] += P^T dO
        }
        umma_commit(&bars[2]);
      }
      mbar_wait(&bars[2], ph_a);      // P consumed: the buffer can take dS
      ph_a ^= 1;
      __syncwarp();
      tc_fence_after();
#pragma unroll
      for (int c8 = 0; c8 < HC / 8; ++c8) {
        const int col = col0 + c8 * 8;
        const int region = col >> 6, c16 = (col & 63) >> 3;
        *reinterpret_cast<uint4*>(sPD + region * MTILE_BYTES + sw_off(r, c16)) =
            make_uint4(dsp[MODE == 0 ? c8 * 4 + 0 : 0], dsp[MODE == 0 ? c8 * 4 + 1 : 0], dsp[MODE == 0 ? c8 * 4 + 2 : 0],
                       dsp[MODE == 0 ? c8 * 4 + 3 : 0]);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < MT / 16; ++kk) {
          const uint64_t aDS = umma_desc_sw128(smem_u32(sPD) + kk * 2048, MTILE_BYTES, 1024);
          const uint64_t bQ = umma_desc_sw128(smem_u32(sQ) + kk * 2048, MTILE_BYTES, 1024);
          umma_f16(tmem + 320, aDS, bQ, id_mm, (it != 0 || kk != 0) ? 1u : 0u);      // dK[k,d] += dS^T Q
        }
        umma_commit(&bars[3]);
      }
      mbar_wait(&bars[3], ph_b);      // Q, dO, dS of this step are consumed
      ph_b ^= 1;
      __syncwarp();
      tc_fence_after();
    } else {
      if (tid == 0) {
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < MT / 16; ++kk) {
          // reduction over keys: dS K-major (two 64-wide regions), K tile MN-major
          const uint64_t aDSk = umma_desc_sw128(smem_u32(sPD) + (kk >> 2) * MTILE_BYTES + (kk & 3) * 32, 16, 1024);
          const uint64_t bK = umma_desc_sw128(smem_u32(sK) + kk * 2048, MTILE_BYTES, 1024);
          umma_f16(tmem + 256, aDSk, bK, id_km, (it != 0 || kk != 0) ? 1u : 0u);     // dQ[q,d] += dS K
        }
        umma_commit(&bars[2]);
      }
      mbar_wait(&bars[2], ph_a);      // K, V, dS of this step are consumed
      ph_a ^= 1;
      __syncwarp();
      tc_fence_after();
    }
  }

  // ---- epilogue: accumulators -> packed [M, 3H] gradient (+ bias-gradient column sums) ----
  {
    const size_t grow = static_cast<size_t>(b) * P.S + outer * MT + r;
    T* base = reinterpret_cast<T*>(P.out) + grow * (3 * P.H) + h * MD + hf * 32;
    constexpr int NACC = MODE == 0 ? 2 : 1;
#pragma unroll
    for (int a = 0; a < NACC; ++a) {
      // MODE 0: a = 0 -> dV (TMEM 256, column block 2), a = 1 -> dK (TMEM 320, column block 1). MODE 1: dQ (block 0)
      const int which = MODE == 0 ? (a == 0 ? 2 : 1) : 0;
      T* dst = base + which * P.H;
      uint32_t rr[32];
      tmem_ld_32x32(t_row + 256 + a * 64 + hf * 32, rr);
      tmem_ld_wait();
      if (outer * MT + r < P.S) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 u;
          u.x = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 0]), __uint_as_float(rr[i * 8 + 1]));
          u.y = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 2]), __uint_as_float(rr[i * 8 + 3]));
          u.z = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 4]), __uint_as_float(rr[i * 8 + 5]));
          u.w = Cvt<T>::pack2(__uint_as_float(rr[i * 8 + 6]), __uint_as_float(rr[i * 8 + 7]));
          reinterpret_cast<uint4*>(dst)[i] = u;
        }
      }
      if (P.dbias != nullptr) {
        float cs[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) cs[i] = __uint_as_float(rr[i]);
        warp_colsum<32>(cs, lane);
        atomicAdd(&sBias[a * MD + hf * 32 + lane], cs[0]);
      }
    }
    tc_fence_before();
    __syncthreads();
    if (P.dbias != nullptr && tid < NACC * MD) {
      const int a = tid / MD, c = tid % MD;
      const int which = MODE == 0 ? (a == 0 ? 2 : 1) : 0;
      atomicAdd(P.dbias + which * P.H + h * MD + c, sBias[tid]);
    }
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

constexpr int MULTI_FWD_SMEM = 3 * MTILE_BYTES + MPBUF_BYTES + (MAX_S + 2 * MT) * 4 + 64 + 1024;
constexpr int MULTI_BWD_SMEM = 4 * MTILE_BYTES + MPBUF_BYTES + (MAX_S + 2 * MD) * 4 + 64 + 1024;

template <typename K>
int prepare_kernel(K kern, int smem, bool* done) {
  if (!*done) {
    HS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    HS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    *done = true;
  }
  return 0;
}

}  // namespace

bool attn_multi_supported(int S, int head_dim) {
  return head_dim == MD && S >= 16 && S % 16 == 0 && S <= MAX_S;
}

// cudaFuncSetAttribute is per device: remember which devices have been prepared for which kernel
template <typename K>
static int prepare_on_device(K kern, int smem, bool (&done)[64]) {
  int dev = 0;
  HS_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return prepare_kernel(kern, smem, &done[63]) ;
  return prepare_kernel(kern, smem, &done[dev]);
}

int attn_multi_fwd(const void* qkv, const float* mask, void* ctx, float* lse, int B, int S, int heads, int H,
                   float dropout_p, const long long* seed, unsigned site, int dtype, cudaStream_t s) {
  CUtensorMap tm;
  if (make_tmap_3d_16b(&tm, qkv, 3ull * H, static_cast<uint64_t>(S), static_cast<uint64_t>(B), 3ull * H,
                       3ull * H * S, 64, MT)) return 1;
  const int nt = (S + MT - 1) / MT;
  const float scale = 1.0f / sqrtf(static_cast<float>(H / heads));      // 1/sqrt(d), bert_modeling.py:361-362
  MultiParams P{mask, ctx, nullptr, nullptr, lse, nullptr, B, heads, H, S, nt, scale, dropout_p, seed, site};
  const int grid = B * heads * nt;
  if (dtype == HS_F16) {
    static bool done[64] = {false};
    if (prepare_on_device(attn_fwd_multi_kernel<__half>, MULTI_FWD_SMEM, done)) return 1;
    (void)launch_k(attn_fwd_multi_kernel<__half>, dim3(grid), dim3(MTHREADS), MULTI_FWD_SMEM, s, tm, P);
  } else if (dtype == HS_BF16) {
    static bool done[64] = {false};
    if (prepare_on_device(attn_fwd_multi_kernel<__nv_bfloat16>, MULTI_FWD_SMEM, done)) return 1;
    (void)launch_k(attn_fwd_multi_kernel<__nv_bfloat16>, dim3(grid), dim3(MTHREADS), MULTI_FWD_SMEM, s, tm, P);
  } else { set_last_error("hs_attention_fwd: dtype must be f16 or bf16"); return 1; }
  return post_launch("attn_fwd_multi_kernel");
}

int attn_multi_bwd(const void* qkv, const float* mask, const void* ctx, const void* dctx, const float* lse,
                   void* dqkv, float* dbias, int B, int S, int heads, int H, float dropout_p, const long long* seed,
                   unsigned site, int dtype, cudaStream_t s) {
  CUtensorMap tmQ, tmD;
  if (make_tmap_3d_16b(&tmQ, qkv, 3ull * H, static_cast<uint64_t>(S), static_cast<uint64_t>(B), 3ull * H,
                       3ull * H * S, 64, MT)) return 1;
  if (make_tmap_3d_16b(&tmD, dctx, static_cast<uint64_t>(H), static_cast<uint64_t>(S), static_cast<uint64_t>(B),
                       static_cast<uint64_t>(H), static_cast<uint64_t>(H) * S, 64, MT)) return 1;
  const int nt = (S + MT - 1) / MT;
  const float scale = 1.0f / sqrtf(static_cast<float>(H / heads));
  MultiParams P{mask, dqkv, ctx, dctx, const_cast<float*>(lse), dbias, B, heads, H, S, nt, scale, dropout_p, seed, site};
  const int grid = B * heads * nt;
#define HS_MULTI_BWD(T)                                                                                  \
  do {                                                                                                   \
    static bool d0[64] = {false}, d1[64] = {false};                                                      \
    if (prepare_on_device(attn_bwd_multi_kernel<T, 0>, MULTI_BWD_SMEM, d0)) return 1;                    \
    if (prepare_on_device(attn_bwd_multi_kernel<T, 1>, MULTI_BWD_SMEM, d1)) return 1;                    \
    (void)launch_k(attn_bwd_multi_kernel<T, 0>, dim3(grid), dim3(MTHREADS), MULTI_BWD_SMEM, s, tmQ, tmD, P);                     \
    if (post_launch("attn_bwd_multi_kernel<dKdV>")) return 1;                                            \
    (void)launch_k(attn_bwd_multi_kernel<T, 1>, dim3(grid), dim3(MTHREADS), MULTI_BWD_SMEM, s, tmQ, tmD, P);                     \
  } while (0)
  if (dtype == HS_F16) HS_MULTI_BWD(__half);
  else if (dtype == HS_BF16) HS_MULTI_BWD(__nv_bfloat16);
  else { set_last_error("hs_attention_bwd: dtype must be f16 or bf16"); return 1; }
#undef HS_MULTI_BWD
  return post_launch("attn_bwd_multi_kernel<dQ>");
}

}  // namespace hs
