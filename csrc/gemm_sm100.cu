// hetseq_b200 -- persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   C[M,N] = epilogue( alpha * sum_k A(m,k) * B(n,k) )        fp16/bf16 in, fp32 accumulate in TMEM
//
// One CTA per SM loops over 128 x BN output tiles. Warp roles:
//   warp 0   TMA producer: cp.async.bulk.tensor tiles of A and B into a STAGES-deep smem ring
//   warp 1   MMA issuer:   one elected thread issues tcgen05.mma (128 x BN x 16) per k-step, commits
//                          ring slots back to the producer and finished accumulators to the epilogue
//   warps 2-9 epilogue:    tcgen05.ld the accumulator (double-buffered in TMEM, so the next tile's
//                          MMAs overlap this tile's epilogue), apply bias / GELU / tanh / dGELU, convert,
//                          stage 32-row x 128-byte boxes in swizzled smem and hand them to the TMA:
//                          plain bulk store, or bulk REDUCE-ADD (fp32) when accumulating -- which is also
//                          what makes split-K work for the wgrad shapes with few output tiles
// Operands may be K-major or MN-major in global memory (both are expressed through the TMA box
// shape and the UMMA shared-memory descriptor), so forward, dgrad and wgrad all run without any
// transpose pass:  forward X*W^T = (K,K);  dgrad dY*W = (K,MN);  wgrad dY^T*X = (MN,MN).
//
// Replaces the cuBLAS calls behind F.linear / autograd in the reference
// (bert_modeling.py:352-354,387-391,409-413,423-427,514-515,525-528,547; SURVEY.md K1,K5-K7,K9-K11,K14)
// and the TorchScript bias_gelu / bias_tanh fusions (bert_modeling.py:104-130,166-172).
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

#include <cstdlib>
#include <cstring>
#include <mutex>

namespace hs {

constexpr int BM = 128;
constexpr int BK = 64;               // 64 x 16-bit = one 128-byte swizzle row
constexpr int A_STAGE_BYTES = BM * 128;
constexpr int EPI_WARPS = 8;             // two warps per TMEM lane quarter, each takes half the columns
constexpr int GEMM_THREADS = 64 + EPI_WARPS * 32;

struct GemmParams {
  void* C;
  const void* bias;
  void* aux_out;
  const void* aux_in;
  float* colsum_out;   // non-NULL: column sums of the produced C tile are atomically added here (bias gradient)
  int M, N, K, ldc, ldaux;
  int tiles_m, tiles_n, num_kb;
  int a_mn, b_mn;
  uint32_t idesc;
  uint32_t a_lbo, a_sbo, a_kadv;  // bytes
  uint32_t b_lbo, b_sbo, b_kadv;
  int act, accumulate;
  float alpha;
  const float* alpha_dev;
  int vec_ok;    // C/aux rows are 16-byte aligned -> vector loads/stores allowed
  int tma_store; // epilogue goes through smem + TMA (needs vec_ok); else per-thread global stores
  int split_k;   // >1: several work items per output tile, each reduce-adds its K range
  int kb_per_split;
};

// CG = CTAs per MMA (tcgen05 cta_group). CG == 2: a CTA pair owns a 256 x BN tile; each CTA stages its own
// 128 rows of A and its own BN/2 rows of B (the MMA reads B from both CTAs' smem), so a stage is 32 KiB
// instead of 48 KiB and the ring is 6 deep -- more bytes in flight per SM for the same shared memory.
// TWO_BOX (the epilogue that stores TWO outputs per chunk, GELU and GELU'): a second staging box per epilogue warp
// so that neither store waits for the other to have been read out of shared memory, paid for with one ring stage.
template <int BN, int CG = 1, bool TWO_BOX = false> struct GemmCfg {
  static constexpr int STAGES = TWO_BOX ? 5 : ((BN == 256 && CG == 1) ? 4 : 6);
  static constexpr int B_STAGE_BYTES = (BN / CG) * 128;
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int BOX_BYTES = 4096;                  // 32 rows x 128 bytes
  static constexpr int STAGING_BYTES = EPI_WARPS * BOX_BYTES * (TWO_BOX ? 2 : 1);
  static constexpr int SMEM_BYTES = STAGES * (A_STAGE_BYTES + B_STAGE_BYTES) + STAGING_BYTES +
                                    (2 * STAGES + 4) * 8 /*mbarriers*/ + 16 /*tmem slot*/ + 1024 /*align*/;
};

// ---- epilogue helpers -------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void load_row32(const T* __restrict__ p, int n0, int N, bool vec, float (&o)[32]) {
  if (vec && n0 + 32 <= N) {
    if constexpr (sizeof(T) == 2) {
      const uint4* q = reinterpret_cast<const uint4*>(p + n0);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 u = q[i];
        float2 f0 = Cvt<T>::unpack2(u.x), f1 = Cvt<T>::unpack2(u.y), f2 = Cvt<T>::unpack2(u.z),
               f3 = Cvt<T>::unpack2(u.w);
        o[i * 8 + 0] = f0.x; o[i * 8 + 1] = f0.y; o[i * 8 + 2] = f1.x; o[i * 8 + 3] = f1.y;
        o[i * 8 + 4] = f2.x; o[i * 8 + 5] = f2.y; o[i * 8 + 6] = f3.x; o[i * 8 + 7] = f3.y;
      }
    } else {
      const float4* q = reinterpret_cast<const float4*>(p + n0);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float4 u = q[i];
        o[i * 4 + 0] = u.x; o[i * 4 + 1] = u.y; o[i * 4 + 2] = u.z; o[i * 4 + 3] = u.w;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = (n0 + i < N) ? Cvt<T>::to_f(p[n0 + i]) : 0.0f;
  }
}

template <typename T>
__device__ __forceinline__ void store_row32(T* __restrict__ p, int n0, int N, bool vec, const float (&v)[32]) {
  if (vec && n0 + 32 <= N) {
    if constexpr (sizeof(T) == 2) {
      uint4* q = reinterpret_cast<uint4*>(p + n0);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 u;
        u.x = Cvt<T>::pack2(v[i * 8 + 0], v[i * 8 + 1]);
        u.y = Cvt<T>::pack2(v[i * 8 + 2], v[i * 8 + 3]);
        u.z = Cvt<T>::pack2(v[i * 8 + 4], v[i * 8 + 5]);
        u.w = Cvt<T>::pack2(v[i * 8 + 6], v[i * 8 + 7]);
        q[i] = u;
      }
    } else {
      float4* q = reinterpret_cast<float4*>(p + n0);
#pragma unroll
      for (int i = 0; i < 8; ++i) q[i] = make_float4(v[i * 4 + 0], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (n0 + i < N) p[n0 + i] = Cvt<T>::from_f(v[i]);
  }
}

// ---- the kernel -------------------------------------------------------------------------------
// ACT >= 0: the activation is a compile-time constant (the kernel only carries that epilogue: the all-in-one
// epilogue is ~8.6 K instructions and ncu showed 19 % of the epilogue warps' stalls as instruction fetch);
// ACT < 0: read it from GemmParams at run time.
template <int BN, typename InT, typename OutT, int CG, int ACT, bool TWO_BOX = false>
__device__ __forceinline__ void gemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                                          const CUtensorMap& tmAux, const GemmParams& p) {
  using Cfg = GemmCfg<BN, CG, TWO_BOX>;
  constexpr int TILE_M = BM * CG;            // rows of C per work item
  const int cta_rank = CG == 2 ? static_cast<int>(cluster_ctarank()) : 0;
  const bool leader = cta_rank == 0;
  const int worker = CG == 2 ? (blockIdx.x >> 1) : blockIdx.x;        // index of this CTA (pair) among the workers
  const int nworkers = CG == 2 ? (gridDim.x >> 1) : gridDim.x;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int B_STAGE_BYTES = Cfg::B_STAGE_BYTES;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_STAGE_BYTES;
  uint8_t* sStage = sB + STAGES * B_STAGE_BYTES;   // 1024-aligned: every stage size is a multiple of 1024
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sStage + Cfg::STAGING_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  // broadcast from lane 0: tells the compiler the role branches below are warp-uniform, so loop state of the
  // producer and MMA warps lives in uniform registers (no R2UR / waterfall code around every TMA and MMA)
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int num_work = num_tiles * p.split_k;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.tma_store) tma_prefetch_desc(&tmC);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tfull_bar[a], 1);
        mbar_init(&tempty_bar[a], EPI_WARPS * CG);  // one arrive per epilogue warp (of both CTAs of a pair)
      }
      fence_mbar_init();
    }
    __syncwarp();
    if constexpr (CG == 2) tmem_alloc_2cta(tmem_slot, Cfg::TMEM_COLS);
    else tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_sync();   // barriers, TMEM and descriptors are set up; operands of the previous kernel are visible from here on

  if (warp == 0) {
    // ===================== TMA producer =====================
    // The whole warp walks the loop (uniform control flow); one elected lane issues the copies.
    // (Measured and rejected: a second cursor that prefetches boxes into L2 ten k-blocks ahead. The layer
    // shapes did not move and 8192^3 dropped from 2050 to 1410 TFLOP/s -- the L2 request path is the limiter,
    // not HBM latency.)
    int stage = 0;
    uint32_t phase = 0;
    const bool a_mn = p.a_mn != 0, b_mn = p.b_mn != 0;
    for (int work = worker; work < num_work; work += nworkers) {
      const int tile = work % num_tiles, ks = work / num_tiles;
      const int tm = tile % p.tiles_m, tn = tile / p.tiles_m;
      const int kb0 = ks * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
      const int m_row = tm * TILE_M + cta_rank * BM;          // this CTA's rows of A
      const int n_row = tn * BN + cta_rank * (BN / CG);       // this CTA's rows of B
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one_sync()) {
          // the leader's barrier collects the bytes of both CTAs of the pair
          if (leader) mbar_expect_tx(&full_bar[stage], CG * (A_STAGE_BYTES + B_STAGE_BYTES));
          uint8_t* a_dst = sA + stage * A_STAGE_BYTES;
          uint8_t* b_dst = sB + stage * B_STAGE_BYTES;
          auto load = [&](void* dst, const CUtensorMap* tm_, int c0, int c1) {
            if constexpr (CG == 2) tma_load_2d_2cta(dst, tm_, &full_bar[stage], c0, c1);
            else tma_load_2d(dst, tm_, &full_bar[stage], c0, c1);
          };
          if (!a_mn) {
            load(a_dst, &tmA, kb * BK, m_row);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) load(a_dst + j * (BK * 128), &tmA, m_row + j * 64, kb * BK);
          }
          if (!b_mn) {
            load(b_dst, &tmB, kb * BK, n_row);
          } else {
#pragma unroll
            for (int j = 0; j < BN / CG / 64; ++j) load(b_dst + j * (BK * 128), &tmB, n_row + j * 64, kb * BK);
          }
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA of a pair only) =====================
    int stage = 0;
    uint32_t phase = 0;
    int as = 0;
    uint32_t aphase = 0;
    const uint64_t adesc0 = umma_desc_sw128(smem_u32(sA), p.a_lbo, p.a_sbo);
    const uint64_t bdesc0 = umma_desc_sw128(smem_u32(sB), p.b_lbo, p.b_sbo);
    const uint32_t a_kadv16 = p.a_kadv >> 4, b_kadv16 = p.b_kadv >> 4;
    if (leader)
    for (int work = worker; work < num_work; work += nworkers) {
      const int ks = work / num_tiles;
      const int kb0 = ks * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
      mbar_wait(&tempty_bar[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one_sync()) {
          // descriptors differ from the stage-0 ones only in the 14-bit start-address field (bytes >> 4)
          const uint64_t adesc = adesc0 + static_cast<uint64_t>(stage * (A_STAGE_BYTES >> 4));
          const uint64_t bdesc = bdesc0 + static_cast<uint64_t>(stage * (B_STAGE_BYTES >> 4));
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            if constexpr (CG == 2)
              umma_f16_2cta(d_tmem, adesc + static_cast<uint64_t>(k * a_kadv16), bdesc + static_cast<uint64_t>(k * b_kadv16),
                            p.idesc, (kb != kb0 || k != 0) ? 1u : 0u);
            else
              umma_f16(d_tmem, adesc + static_cast<uint64_t>(k * a_kadv16), bdesc + static_cast<uint64_t>(k * b_kadv16),
                       p.idesc, (kb != kb0 || k != 0) ? 1u : 0u);
          }
          if constexpr (CG == 2) {
            umma_commit_2cta(&empty_bar[stage], 3);                  // frees the slot in both CTAs
            if (kb == kb1 - 1) umma_commit_2cta(&tfull_bar[as], 3);   // accumulator complete (both CTAs)
          } else {
            umma_commit(&empty_bar[stage]);                  // ring slot free once these MMAs retire
            if (kb == kb1 - 1) umma_commit(&tfull_bar[as]);   // accumulator complete
          }
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    const int q = warp & 3;            // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;  // which half of the tile's columns this warp drains
    const InT* bias = reinterpret_cast<const InT*>(p.bias);
    OutT* C = reinterpret_cast<OutT*>(p.C);
    const bool vec = p.vec_ok != 0;
    const bool tma = p.tma_store != 0;
    const float alpha = p.alpha * (p.alpha_dev != nullptr ? *p.alpha_dev : 1.0f);
    constexpr int CW = (sizeof(OutT) == 2) ? 64 : 32;   // columns per 128-byte staging row
    constexpr int NCH = (BN / 2) / CW;                  // chunks per tile for this warp
    const int act = ACT >= 0 ? ACT : p.act;
    const bool use_aux = act == HS_ACT_DGELU_MUL || act == HS_ACT_ADD_AUX || act == HS_ACT_MUL_AUX;
    // TWO_BOX: the derivative goes through the warp's first box, the output through its second one; each store
    // then only has to wait for the store of the PREVIOUS chunk from the same box (at most one younger group pending)
    uint8_t* stg_aux = sStage + (warp - 2) * (Cfg::BOX_BYTES * (TWO_BOX ? 2 : 1));
    uint8_t* stg = stg_aux + (TWO_BOX ? Cfg::BOX_BYTES : 0);
    // The epilogue operand (pre-activation for dGELU, residual-path gradient for ADD) is this thread's own row:
    // CW 16-bit values = CW/8 16-byte loads. ncu showed the warps stalled on the long scoreboard right after
    // these loads, so they are issued early: an L2 prefetch of the warp's whole part of the tile and the loads
    // of the first chunk go out BEFORE the wait for the accumulator (the main loop hides HBM latency), the
    // loads of the next chunk right after the current chunk's values have been consumed (the output staging
    // and the next TMEM read hide L2 latency). One register buffer, so no extra pressure.
    uint4 ax[CW / 8];
    bool ax_ok = false;
    auto aux_issue = [&](int m_, int n0_) {
      ax_ok = use_aux && vec && m_ < p.M && n0_ + CW <= p.N;
      if (ax_ok) {
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const InT*>(p.aux_in) +
                                                          static_cast<size_t>(m_) * p.ldaux + n0_);
#pragma unroll
        for (int j = 0; j < CW / 8; ++j) ax[j] = __ldg(src + j);
      }
    };
    int as = 0;
    uint32_t aphase = 0;
    for (int work = worker; work < num_work; work += nworkers) {
      const int tile = work % num_tiles, ks = work / num_tiles;
      const int tm = tile % p.tiles_m, tn = tile / p.tiles_m;
      const int n_base = tn * BN;
      const int m0 = tm * TILE_M + cta_rank * BM + q * 32;
      const int m = m0 + lane;
      const bool row_ok = m < p.M;
      if (use_aux && row_ok) {
        const InT* arow = reinterpret_cast<const InT*>(p.aux_in) + static_cast<size_t>(m) * p.ldaux;
#pragma unroll
        for (int cc = 1; cc < NCH; ++cc) {
          const int n0 = n_base + half * (BN / 2) + cc * CW;
          if (n0 < p.N) asm volatile("prefetch.global.L2 [%0];" ::"l"(arow + n0));
        }
      }
      aux_issue(m, n_base + half * (BN / 2));
      mbar_wait(&tfull_bar[as], aphase);
      __syncwarp();
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
      const bool add_bias = bias != nullptr && ks == 0;
#pragma unroll 1
      for (int cc = 0; cc < NCH; ++cc) {
        const int c0 = half * (BN / 2) + cc * CW;
        const int n0 = n_base + c0;
        if (n0 >= p.N) break;  // warp-uniform
        float v[CW];
        const bool bias_vec = add_bias && n0 + CW <= p.N;
#pragma unroll
        for (int sub = 0; sub < CW / 32; ++sub) {
          uint32_t r[32];
          tmem_ld_32x32(t_row + c0 + sub * 32, r);
          if (bias_vec) {
            // bias loads go out under the TMEM read; alpha and the bias are one FMA per element
            uint4 bu[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) bu[j] = *reinterpret_cast<const uint4*>(bias + n0 + sub * 32 + j * 8);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 f0 = Cvt<InT>::unpack2(bu[j].x), f1 = Cvt<InT>::unpack2(bu[j].y), f2 = Cvt<InT>::unpack2(bu[j].z),
                           f3 = Cvt<InT>::unpack2(bu[j].w);
              const float bf[8] = {f0.x, f0.y, f1.x, f1.y, f2.x, f2.y, f3.x, f3.y};
#pragma unroll
              for (int i = 0; i < 8; ++i)
                v[sub * 32 + j * 8 + i] = fmaf(__uint_as_float(r[j * 8 + i]), alpha, bf[i]);
            }
          } else {
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[sub * 32 + i] = __uint_as_float(r[i]) * alpha;
          }
        }
        if (add_bias && !bias_vec) {
#pragma unroll
          for (int i = 0; i < CW; ++i)
            if (n0 + i < p.N) v[i] += Cvt<InT>::to_f(bias[n0 + i]);
        }
        // ---- activation (pre-activation goes out first when requested)
        if (act == HS_ACT_GELU || act == HS_ACT_GELU_DAUX) {
          const bool daux = act == HS_ACT_GELU_DAUX;
          float av[CW];     // what goes to aux_out: the pre-activation, or the derivative of the activation
          if (daux) {
#pragma unroll
            for (int i = 0; i < CW; ++i) gelu_and_grad(v[i], v[i], av[i]);
          } else {
#pragma unroll
            for (int i = 0; i < CW; ++i) av[i] = v[i];
          }
          if (p.aux_out != nullptr) {
            if constexpr (sizeof(OutT) == 2) {
              if (tma) {
                if (lane == 0) {
                  if constexpr (TWO_BOX) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                  else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                }
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  uint4 u;
                  u.x = Cvt<InT>::pack2(av[j * 8 + 0], av[j * 8 + 1]); u.y = Cvt<InT>::pack2(av[j * 8 + 2], av[j * 8 + 3]);
                  u.z = Cvt<InT>::pack2(av[j * 8 + 4], av[j * 8 + 5]); u.w = Cvt<InT>::pack2(av[j * 8 + 6], av[j * 8 + 7]);
                  *reinterpret_cast<uint4*>(stg_aux + lane * 128 + ((j ^ (lane & 7)) << 4)) = u;
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                               ::"l"(reinterpret_cast<uint64_t>(&tmAux)), "r"(smem_u32(stg_aux)), "r"(n0), "r"(m0) : "memory");
                  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
              } else if (row_ok) {
#pragma unroll
                for (int sub = 0; sub < CW / 32; ++sub) {
                  float t32[32];
#pragma unroll
                  for (int i = 0; i < 32; ++i) t32[i] = av[sub * 32 + i];
                  store_row32<InT>(reinterpret_cast<InT*>(p.aux_out) + static_cast<size_t>(m) * p.ldaux, n0 + sub * 32, p.N, vec, t32);
                }
              }
            }
          }
          if (act == HS_ACT_GELU) {
#pragma unroll
            for (int i = 0; i < CW; ++i) v[i] = gelu_ref(v[i]);
          }
        } else if (act == HS_ACT_TANH) {
#pragma unroll
          for (int i = 0; i < CW; ++i) v[i] = tanhf(v[i]);
        } else if (use_aux) {
          if (ax_ok) {
#pragma unroll
            for (int j = 0; j < CW / 8; ++j) {
              const float2 f0 = Cvt<InT>::unpack2(ax[j].x), f1 = Cvt<InT>::unpack2(ax[j].y),
                           f2 = Cvt<InT>::unpack2(ax[j].z), f3 = Cvt<InT>::unpack2(ax[j].w);
              const float x[8] = {f0.x, f0.y, f1.x, f1.y, f2.x, f2.y, f3.x, f3.y};
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                if (act == HS_ACT_DGELU_MUL) v[j * 8 + i] *= gelu_grad_ref(x[i]);
                else if (act == HS_ACT_MUL_AUX) v[j * 8 + i] *= x[i];
                else v[j * 8 + i] += x[i];
              }
            }
          } else if (row_ok) {   // ragged edge / unaligned operand: synchronous element-wise loads
#pragma unroll
            for (int sub = 0; sub < CW / 32; ++sub) {
              float x[32];
              load_row32<InT>(reinterpret_cast<const InT*>(p.aux_in) + static_cast<size_t>(m) * p.ldaux, n0 + sub * 32, p.N, false, x);
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                if (act == HS_ACT_DGELU_MUL) v[sub * 32 + i] *= gelu_grad_ref(x[i]);
                else if (act == HS_ACT_MUL_AUX) v[sub * 32 + i] *= x[i];
                else v[sub * 32 + i] += x[i];
              }
            }
          }
          if (cc + 1 < NCH) aux_issue(m, n0 + CW);
        }
        // ---- output
        if (tma) {
          if (lane == 0) {   // the staging box is free again
            if constexpr (TWO_BOX) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          }
          __syncwarp();
          if constexpr (sizeof(OutT) == 2) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              uint4 u;
              u.x = Cvt<OutT>::pack2(v[j * 8 + 0], v[j * 8 + 1]); u.y = Cvt<OutT>::pack2(v[j * 8 + 2], v[j * 8 + 3]);
              u.z = Cvt<OutT>::pack2(v[j * 8 + 4], v[j * 8 + 5]); u.w = Cvt<OutT>::pack2(v[j * 8 + 6], v[j * 8 + 7]);
              *reinterpret_cast<uint4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = u;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *reinterpret_cast<float4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                  make_float4(v[j * 4 + 0], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (p.accumulate) {
              asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                           ::"l"(reinterpret_cast<uint64_t>(&tmC)), "r"(smem_u32(stg)), "r"(n0), "r"(m0) : "memory");
            } else {
              asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                           ::"l"(reinterpret_cast<uint64_t>(&tmC)), "r"(smem_u32(stg)), "r"(n0), "r"(m0) : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        } else if (row_ok) {
          OutT* crow = C + static_cast<size_t>(m) * p.ldc;
#pragma unroll
          for (int sub = 0; sub < CW / 32; ++sub) {
            float t32[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) t32[i] = v[sub * 32 + i];
            if (p.accumulate) {
              float o[32];
              load_row32<OutT>(crow, n0 + sub * 32, p.N, vec, o);
#pragma unroll
              for (int i = 0; i < 32; ++i) t32[i] += o[i];
            }
            store_row32<OutT>(crow, n0 + sub * 32, p.N, vec, t32);
          }
        }
        // ---- bias gradient of the Linear this tile is the output gradient of: column sums over the warp's
        // 32 rows by a halving butterfly, then one RED per column
        if (p.colsum_out != nullptr) {
          if (!row_ok) {
#pragma unroll
            for (int i = 0; i < CW; ++i) v[i] = 0.f;
          }
          warp_colsum<CW>(v, lane);
#pragma unroll
          for (int i = 0; i < CW / 32; ++i) {
            const int n = n0 + lane * (CW / 32) + i;
            if (n < p.N) atomicAdd(p.colsum_out + n, v[i]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 2) mbar_arrive_leader(&tempty_bar[as]);   // the pair's MMA issuer lives in CTA 0
        else mbar_arrive(&tempty_bar[as]);
      }
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    // every bulk store of this warp must have completed before the CTA (and its smem) goes away
    if (tma && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all();   // the leader's MMAs read the peer's smem: nobody leaves early
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if constexpr (CG == 2) tmem_dealloc_2cta(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

template <int BN, typename InT, typename OutT>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmAux,
                    const GemmParams p) {
  gemm_body<BN, InT, OutT, 1, -1>(tmA, tmB, tmC, tmAux, p);
}

// CTA-pair variant: tcgen05.mma.cta_group::2 with UMMA_M = 256 (the pair's 256 x 256 tile); one instance per
// epilogue kind (ACT = -1: generic)
template <typename InT, typename OutT, int ACT, bool TWO_BOX = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmAux,
                         const GemmParams p) {
  gemm_body<256, InT, OutT, 2, ACT, TWO_BOX>(tmA, tmB, tmC, tmAux, p);
}

// ---- host side --------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

static int make_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
                        uint32_t box_inner, uint32_t box_outer, int elem_bytes);

int make_tmap_2d_16b(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
                     uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_2d(out, base, inner, outer, ld, box_inner, box_outer, 2);
}

static int make_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
                        uint32_t box_inner, uint32_t box_outer, int elem_bytes) {
  // cuTensorMapEncodeTiled is a driver-API call: it needs the primary context to be current on THIS
  // thread. A thread that has only ever inherited the default device (e.g. PyTorch's autograd worker
  // on device 0) has no context bound until its first context-requiring runtime call -- force that.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    (void)cudaFree(nullptr);
    ctx_bound = true;
  }
  PFN_encodeTiled enc = get_encode_fn();
  if (enc == nullptr) {
    set_last_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    return 1;
  }
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * static_cast<uint64_t>(elem_bytes)};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT16, 2,
                   const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed (%d): base=%p inner=%llu outer=%llu ld=%llu box=%ux%u", (int)r,
                   base, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld, box_inner,
                   box_outer);
    return 1;
  }
  return 0;
}

// 3-D tiled map (128-byte swizzle) over a 16-bit tensor [outer][mid][inner] (inner contiguous): a box of
// box_inner x box_mid x 1 elements lands in shared memory exactly like the 2-D box of make_tmap_2d_16b, but rows
// beyond `mid` are out of bounds -- zero-filled -- instead of running into the next `outer` slice. The attention
// kernels use it to read a sequence's last, partial 128-row tile.
int make_tmap_3d_16b(CUtensorMap* out, const void* base, uint64_t inner, uint64_t mid, uint64_t outer,
                     uint64_t ld_mid, uint64_t ld_outer, uint32_t box_inner, uint32_t box_mid) {
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    (void)cudaFree(nullptr);
    ctx_bound = true;
  }
  PFN_encodeTiled enc = get_encode_fn();
  if (enc == nullptr) {
    set_last_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    return 1;
  }
  cuuint64_t dims[3] = {inner, mid, outer};
  cuuint64_t strides[2] = {ld_mid * 2ull, ld_outer * 2ull};
  cuuint32_t box[3] = {box_inner, box_mid, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (3-D) failed (%d): base=%p dims=%llu x %llu x %llu", (int)r, base,
                   (unsigned long long)inner, (unsigned long long)mid, (unsigned long long)outer);
    return 1;
  }
  return 0;
}

template <int BN, typename InT, typename OutT, int CG>
static int launch_gemm(const hs_gemm_args& a, cudaStream_t stream) {
  constexpr int TILE_M = BM * CG;
  void (*kern)(CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, GemmParams);
  int variant = 0;
  int smem_bytes = GemmCfg<BN, CG>::SMEM_BYTES;
  // HETSEQ_GEMM_TWO_BOX=0: the GELU + GELU' epilogue with one staging box and a 6-deep ring (comparison)
  static const bool two_box = [] { const char* e = getenv("HETSEQ_GEMM_TWO_BOX"); return e == nullptr || atoi(e) != 0; }();
  if constexpr (CG == 2) {
    kern = gemm_tcgen05_2cta_kernel<InT, OutT, -1>;
    if (a.act == HS_ACT_NONE) { kern = gemm_tcgen05_2cta_kernel<InT, OutT, HS_ACT_NONE>; variant = 1; }
    if constexpr (sizeof(OutT) == 2) {
      if (a.act == HS_ACT_GELU) { kern = gemm_tcgen05_2cta_kernel<InT, OutT, HS_ACT_GELU>; variant = 2; }
      else if (a.act == HS_ACT_DGELU_MUL) { kern = gemm_tcgen05_2cta_kernel<InT, OutT, HS_ACT_DGELU_MUL>; variant = 3; }
      else if (a.act == HS_ACT_ADD_AUX) { kern = gemm_tcgen05_2cta_kernel<InT, OutT, HS_ACT_ADD_AUX>; variant = 4; }
      else if (a.act == HS_ACT_GELU_DAUX && !two_box) { kern = gemm_tcgen05_2cta_kernel<InT, OutT, HS_ACT_GELU_DAUX>; variant = 5; }
      else if (a.act == HS_ACT_GELU_DAUX) {
        kern = gemm_tcgen05_2cta_kernel<InT, OutT, HS_ACT_GELU_DAUX, true>;
        variant = 7;
        smem_bytes = GemmCfg<BN, CG, true>::SMEM_BYTES;
      }
      else if (a.act == HS_ACT_MUL_AUX) { kern = gemm_tcgen05_2cta_kernel<InT, OutT, HS_ACT_MUL_AUX>; variant = 6; }
    }
  } else {
    kern = gemm_tcgen05_kernel<BN, InT, OutT>;
  }
  static bool attr_done[64][8] = {};
  int dev = 0;
  HS_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 64 && !attr_done[dev][variant]) {
    HS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    attr_done[dev][variant] = true;
  }
  CUtensorMap tmA, tmB;
  int rc;
  if (!a.a_mn_major) rc = make_tmap_2d_16b(&tmA, a.A, a.K, a.M, a.lda, BK, BM);
  else               rc = make_tmap_2d_16b(&tmA, a.A, a.M, a.K, a.lda, 64, BK);
  if (rc) return rc;
  if (!a.b_mn_major) rc = make_tmap_2d_16b(&tmB, a.B, a.K, a.N, a.ldb, BK, BN / CG);
  else               rc = make_tmap_2d_16b(&tmB, a.B, a.N, a.K, a.ldb, 64, BK);
  if (rc) return rc;

  GemmParams p;
  p.C = a.C; p.bias = a.bias; p.aux_out = a.aux_out; p.aux_in = a.aux_in; p.colsum_out = a.colsum_out;
  p.M = a.M; p.N = a.N; p.K = a.K; p.ldc = a.ldc; p.ldaux = a.ldaux;
  p.tiles_m = (a.M + TILE_M - 1) / TILE_M;
  p.tiles_n = (a.N + BN - 1) / BN;
  p.num_kb = (a.K + BK - 1) / BK;
  p.a_mn = a.a_mn_major; p.b_mn = a.b_mn_major;
  p.idesc = umma_idesc_f16(a.in_dtype == HS_BF16 ? 1 : 0, a.a_mn_major, a.b_mn_major, TILE_M, BN);
  // K-major, 128B swizzle: 8-row groups 1024 B apart; a k-step of 16 elements is 32 B inside the row.
  // MN-major, 128B swizzle: 64-element MN groups are separate TMA boxes (BK*128 B apart), 8-k groups
  // 1024 B apart; a k-step of 16 rows is 2048 B.
  p.a_lbo = a.a_mn_major ? BK * 128 : 16; p.a_sbo = 1024; p.a_kadv = a.a_mn_major ? 2048 : 32;
  p.b_lbo = a.b_mn_major ? BK * 128 : 16; p.b_sbo = 1024; p.b_kadv = a.b_mn_major ? 2048 : 32;
  p.act = a.act; p.accumulate = a.accumulate; p.alpha = a.alpha; p.alpha_dev = a.alpha_dev;
  const size_t osz = sizeof(OutT);
  bool vec = (reinterpret_cast<uintptr_t>(a.C) % 16 == 0) && ((a.ldc * osz) % 16 == 0);
  if (a.aux_out) vec = vec && (reinterpret_cast<uintptr_t>(a.aux_out) % 16 == 0) && ((a.ldaux * 2) % 16 == 0);
  if (a.aux_in) vec = vec && (reinterpret_cast<uintptr_t>(a.aux_in) % 16 == 0) && ((a.ldaux * 2) % 16 == 0);
  p.vec_ok = vec ? 1 : 0;
  // 16-bit read-modify-write accumulation has no TMA reduce type here: it keeps the per-thread path
  p.tma_store = (vec && !(a.accumulate && sizeof(OutT) == 2)) ? 1 : 0;

  const int tiles = p.tiles_m * p.tiles_n;
  const int sms = num_sms() / CG;   // workers: SMs, or SM pairs
  // split-K: only with the fp32 reduce-add epilogue and no per-column work. When the caller asked for
  // overwrite semantics, C is cleared on the stream first and every split reduce-adds into it.
  p.split_k = 1;
  if (vec && sizeof(OutT) == 4 && a.bias == nullptr && a.act == HS_ACT_NONE && a.colsum_out == nullptr && tiles < sms) {
    double best = 1e30;
    for (int s = 1; s <= 8 && s * 4 <= p.num_kb; ++s) {
      const int items = tiles * s;
      const int waves = (items + sms - 1) / sms;
      const double cost = waves * ((p.num_kb + s - 1) / s + 6.0);   // +6 k-blocks ~ prologue + epilogue of an item
      if (cost < best - 1e-9) { best = cost; p.split_k = s; }
    }
    if (p.split_k > 1 && !a.accumulate) {
      HS_CUDA_OK(cudaMemset2DAsync(a.C, static_cast<size_t>(a.ldc) * sizeof(OutT), 0, static_cast<size_t>(a.N) * sizeof(OutT),
                                   static_cast<size_t>(a.M), stream));
      p.accumulate = 1;
      p.tma_store = 1;
    }
  }
  p.kb_per_split = (p.num_kb + p.split_k - 1) / p.split_k;

  CUtensorMap tmC, tmAux;
  memset(&tmC, 0, sizeof(tmC));
  memset(&tmAux, 0, sizeof(tmAux));
  if (p.tma_store) {
    constexpr int CW = sizeof(OutT) == 2 ? 64 : 32;
    if (make_tmap_2d(&tmC, a.C, a.N, a.M, a.ldc, CW, 32, sizeof(OutT))) return 1;
    if (a.aux_out != nullptr && make_tmap_2d(&tmAux, a.aux_out, a.N, a.M, a.ldaux, 64, 32, 2)) return 1;
  }
  if (a.colsum_out != nullptr && !a.colsum_accumulate)
    HS_CUDA_OK(cudaMemsetAsync(a.colsum_out, 0, sizeof(float) * a.N, stream));
  const int work = tiles * p.split_k;
  const int grid = (work < sms ? work : sms) * CG;
  (void)launch_k(kern, dim3(grid), dim3(GEMM_THREADS), smem_bytes, stream, tmA, tmB, tmC, tmAux, p);
  return post_launch("gemm_tcgen05_kernel");
}

template <int BN, int CG>
static int dispatch_types(const hs_gemm_args& a, cudaStream_t s) {
  if (a.in_dtype == HS_F16) {
    if (a.out_dtype == HS_F16) return launch_gemm<BN, __half, __half, CG>(a, s);
    if (a.out_dtype == HS_F32) return launch_gemm<BN, __half, float, CG>(a, s);
  } else if (a.in_dtype == HS_BF16) {
    if (a.out_dtype == HS_BF16) return launch_gemm<BN, __nv_bfloat16, __nv_bfloat16, CG>(a, s);
    if (a.out_dtype == HS_F32) return launch_gemm<BN, __nv_bfloat16, float, CG>(a, s);
  }
  set_last_error("hs_gemm: unsupported dtype combination in=%d out=%d", a.in_dtype, a.out_dtype);
  return 1;
}

}  // namespace hs

extern "C" int hs_gemm(const hs_gemm_args* args, hs_stream_t stream) {
  using namespace hs;
  if (args == nullptr) { set_last_error("hs_gemm: null args"); return 1; }
  const hs_gemm_args& a = *args;
  if (a.M <= 0 || a.N <= 0 || a.K <= 0) { set_last_error("hs_gemm: empty problem %dx%dx%d", a.M, a.N, a.K); return 1; }
  if (a.lda % 8 || a.ldb % 8) { set_last_error("hs_gemm: lda/ldb must be multiples of 8 (got %d, %d)", a.lda, a.ldb); return 1; }
  if ((reinterpret_cast<uintptr_t>(a.A) | reinterpret_cast<uintptr_t>(a.B)) % 16) {
    set_last_error("hs_gemm: A and B must be 16-byte aligned");
    return 1;
  }
  if ((a.act == HS_ACT_DGELU_MUL || a.act == HS_ACT_ADD_AUX || a.act == HS_ACT_MUL_AUX) && a.aux_in == nullptr) {
    set_last_error("hs_gemm: DGELU_MUL / ADD_AUX / MUL_AUX need aux_in");
    return 1;
  }
  if (a.act == HS_ACT_GELU_DAUX && (a.aux_out == nullptr || a.out_dtype == HS_F32)) {
    set_last_error("hs_gemm: GELU_DAUX needs aux_out and a 16-bit output");
    return 1;
  }
  // Tile-width heuristic: prefer 128x256 tiles (fewer B re-reads) unless 128x128 fills the SMs better.
  const int sms = num_sms();
  const long tm = (a.M + BM - 1) / BM;
  const long t256 = tm * ((a.N + 255) / 256), t128 = tm * ((a.N + 127) / 128);
  auto waves_cost = [&](long tiles, int bn) {  // time ~ waves * tile cost
    const long waves = (tiles + sms - 1) / sms;
    return static_cast<double>(waves) * bn;
  };
  const bool use256 = a.N > 128 && waves_cost(t256, 256) <= waves_cost(t128, 128);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // CTA pairs (cta_group::2, 256 x 256 tiles) whenever the problem has at least one full pair tile;
  // HETSEQ_GEMM_CTA_GROUP=1 forces the single-CTA kernel (debugging / comparison)
  static const int force_cg = [] { const char* e = getenv("HETSEQ_GEMM_CTA_GROUP"); return e ? atoi(e) : 0; }();
  const bool pair = force_cg == 2 || (force_cg != 1 && a.M >= 256 && a.N > 128);
  if (pair) return dispatch_types<256, 2>(a, s);
  return use256 ? dispatch_types<256, 1>(a, s) : dispatch_types<128, 1>(a, s);
}
