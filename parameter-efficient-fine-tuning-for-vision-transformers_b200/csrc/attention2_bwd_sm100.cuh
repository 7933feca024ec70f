// tcgen05 / TMEM / TMA multi-head attention backward (L <= 288, head_dim 64): two kernels, both of the shape
//   two SS products into TMEM -> per-row elementwise softmax-backward in registers -> bf16 operands written back IN PLACE
//   -> TS products (A from TMEM, B = MN-major smem tile) -> scaled bf16 result through a swizzled smem tile and 3-D TMA.
//
//   dQ kernel  (TMEM lane = query row):  S = Q_t K^T, dP = dO_t V^T, dS = P o (dP - delta), dQ_t = scale dS K
//                                        also emits delta = rowsum(dO o O) (O tile comes in by TMA)
//   dKV kernel (TMEM lane = key row):    S^T = K_j Q^T, dP^T = V_j dO^T, P^T, dS^T, dV_j = P^T dO, dK_j = scale dS^T Q
//
// replaces: the autograd backward of FS/kadaptation.py:223-235 / FS/model.py:990-1002.  The keys
// (dQ kernel) / queries (dKV kernel) are walked in NCH chunks of NKC columns so that S and dP fit the 512 TMEM columns:
// (64,1) for L <= 64, (208,1) for L <= 208, (144,2) for L <= 288.  The 8 compute warps split the columns of a chunk:
// warpgroup A takes [0, kCA) ascending and packs its bf16 results at the bottom of its range, warpgroup B takes
// [kCA, NKC) descending and packs at the top, so no thread ever overwrites a column another thread still has to read.
#pragma once
#include "attention2_sm100.cuh"

namespace peft {

template <int NKC, int NCH>
struct Attn2BwdCfg {
  static_assert(NKC % 16 == 0 && NKC >= 64 && NKC <= 208, "chunk width");
  static constexpr int kNKT = NKC * NCH;                               // rows of the per-head smem operands
  static constexpr int kCA = NKC - 32 * ((NKC / 2) / 32);              // warpgroup A: [0, kCA), B: [kCA, NKC)
  static constexpr int kPB = NKC - (NKC - kCA) / 2;                    // first packed column of warpgroup B
  static constexpr int kHeadBytes = kNKT * 128;                        // one per-head operand matrix
  static constexpr int kTileBytes = 128 * 128;
  // ---- dQ kernel
  static constexpr int kDqTmemUsed = 2 * NKC + 64;
  static constexpr int kDqTmemCols = kDqTmemUsed <= 256 ? 256 : 512;
  static constexpr int kDqFixed = 2 * 3 * kTileBytes + kTileBytes + 512 + 1024;   // Q,dO,O x2 | staging | barriers | align
  static constexpr int kDqKVBufs = (4 * kHeadBytes + kDqFixed <= 232448) ? 2 : 1;
  static constexpr int kDqSmemBytes = 2 * kDqKVBufs * kHeadBytes + kDqFixed;
  // ---- dKV kernel
  static constexpr bool kSeparateOut = (2 * NKC + 128 <= 512);
  static constexpr int kOA = ((kCA / 2 + 31) / 32) * 32;               // alias offset inside the S^T / dP^T regions
  static_assert(kSeparateOut || (NCH == 1 && kOA + 64 <= kPB), "dK/dV accumulators do not fit");
  static constexpr int kDvCol = kSeparateOut ? 2 * NKC : kOA;
  static constexpr int kDkCol = kSeparateOut ? 2 * NKC + 64 : NKC + kOA;
  static constexpr int kDkvTmemUsed = kSeparateOut ? 2 * NKC + 128 : 2 * NKC;
  static constexpr int kDkvTmemCols = kDkvTmemUsed <= 256 ? 256 : 512;
  static constexpr int kDkvFixed = 2 * 2 * kTileBytes + 2 * kTileBytes + 2 * kNKT * 4 + 512 + 1024;  // K_j,V_j x2 | staging x2 | lse,delta
  static constexpr int kDkvQDBufs = (4 * kHeadBytes + kDkvFixed <= 232448) ? 2 : 1;
  static constexpr int kDkvSmemBytes = 2 * kDkvQDBufs * kHeadBytes + kDkvFixed;
};

// packed (bf16x2) TMEM column, relative to the region base, that holds chunk columns [16 j, 16 j + 16)
template <int NKC, int NCH>
__device__ __forceinline__ uint32_t bwd_pcol(int j) {
  using C = Attn2BwdCfg<NKC, NCH>;
  const int key = 16 * j;
  return key < C::kCA ? (uint32_t)(key >> 1) : (uint32_t)(NKC - ((NKC - key) >> 1));
}

// One chunk of W columns of the dQ kernel: dS = exp2(S c - lse) (dP - delta), columns >= `valid` masked to 0.
template <int W>
__device__ __forceinline__ void dq_chunk(uint32_t ts, uint32_t tdp, uint32_t tout, int valid, float c, float lse, float delta) {
  uint32_t s[W], d[W], pk[W / 2];
  tm_ld<W>(ts, s);
  tm_ld<W>(tdp, d);
  ptx::tmem_ld_wait();
  if (valid >= W) {
#pragma unroll
    for (int j = 0; j < W; j += 2) {
      const float p0 = ptx::ex2_approx(fmaf(__uint_as_float(s[j]), c, -lse));
      const float p1 = ptx::ex2_approx(fmaf(__uint_as_float(s[j + 1]), c, -lse));
      pk[j >> 1] = ptx::pack_bf16x2(p0 * (__uint_as_float(d[j]) - delta), p1 * (__uint_as_float(d[j + 1]) - delta));
    }
  } else {
#pragma unroll
    for (int j = 0; j < W; j += 2) {
      const float p0 = (j < valid) ? ptx::ex2_approx(fmaf(__uint_as_float(s[j]), c, -lse)) : 0.f;
      const float p1 = (j + 1 < valid) ? ptx::ex2_approx(fmaf(__uint_as_float(s[j + 1]), c, -lse)) : 0.f;
      pk[j >> 1] = ptx::pack_bf16x2(p0 * (__uint_as_float(d[j]) - delta), p1 * (__uint_as_float(d[j + 1]) - delta));
    }
  }
  tm_st_packed<W>(tout, pk);
}

// One chunk of W columns of the dKV kernel (columns = queries): P^T and dS^T, with per-column lse / delta from smem.
template <int W>
__device__ __forceinline__ void dkv_chunk(uint32_t ts, uint32_t tdp, uint32_t tout_p, uint32_t tout_ds, int valid, float c,
                                          const float* __restrict__ lse, const float* __restrict__ delta) {
  uint32_t s[W], d[W], pp[W / 2], pd[W / 2];
  tm_ld<W>(ts, s);
  tm_ld<W>(tdp, d);
  ptx::tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < W; j += 4) {
    const float4 l4 = *reinterpret_cast<const float4*>(lse + j);
    const float4 d4 = *reinterpret_cast<const float4*>(delta + j);
    const float lv[4] = {l4.x, l4.y, l4.z, l4.w};
    const float dv[4] = {d4.x, d4.y, d4.z, d4.w};
    float pv[4], sv[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float pr = (j + q < valid) ? ptx::ex2_approx(fmaf(__uint_as_float(s[j + q]), c, -lv[q])) : 0.f;
      pv[q] = pr;
      sv[q] = pr * (__uint_as_float(d[j + q]) - dv[q]);
    }
    pp[(j >> 1)] = ptx::pack_bf16x2(pv[0], pv[1]);
    pp[(j >> 1) + 1] = ptx::pack_bf16x2(pv[2], pv[3]);
    pd[(j >> 1)] = ptx::pack_bf16x2(sv[0], sv[1]);
    pd[(j >> 1) + 1] = ptx::pack_bf16x2(sv[2], sv[3]);
  }
  tm_st_packed<W>(tout_p, pp);
  tm_st_packed<W>(tout_ds, pd);
}

// =====================================================================================================================
// dQ kernel
// =====================================================================================================================
template <int NKC, int NCH>
__global__ void __launch_bounds__(kA2Threads, 1)
attn2_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                    const __grid_constant__ CUtensorMap tmap_do, const __grid_constant__ CUtensorMap tmap_o,
                    const __grid_constant__ CUtensorMap tmap_out, const Attn2Params p) {
  using C = Attn2BwdCfg<NKC, NCH>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sKV = smem;                                          // [kDqKVBufs][K | V]
  uint8_t* sT = sKV + 2 * C::kDqKVBufs * C::kHeadBytes;         // [2 tiles][Q | dO | O]
  uint8_t* sOut = sT + 6 * C::kTileBytes;                       // dQ staging
  uint64_t* bars = reinterpret_cast<uint64_t*>(sOut + C::kTileBytes);
  uint64_t* kv_full = bars;            // [2]
  uint64_t* kv_empty = bars + 2;       // [2]
  uint64_t* t_full = bars + 4;         // [2]
  uint64_t* t_empty = bars + 6;        // [2]
  uint64_t* sdp_ready = bars + 8;
  uint64_t* ds_ready = bars + 9;
  uint64_t* dq_ready = bars + 10;
  uint64_t* dq_free = bars + 11;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int L = p.L, H = p.H;
  const int TQ = (L + 127) / 128;
  const int total_heads = p.B * H;
  const int my_heads = (total_heads - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int G = my_heads * TQ;
  const int D = H * 64;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_q); ptx::prefetch_tmap(&tmap_kv); ptx::prefetch_tmap(&tmap_do);
    ptx::prefetch_tmap(&tmap_o); ptx::prefetch_tmap(&tmap_out);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&kv_full[i], 1);
      ptx::mbar_init(&kv_empty[i], 1);
      ptx::mbar_init(&t_full[i], 1);
      ptx::mbar_init(&t_empty[i], 1);
    }
    ptx::mbar_init(sdp_ready, 1);
    ptx::mbar_init(ds_ready, 256);
    ptx::mbar_init(dq_ready, 1);
    ptx::mbar_init(dq_free, 256);
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc<C::kDqTmemCols>(tmem_slot);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                            // PDL (common.cuh): no global-memory access above this line

  if (warp == 0) {
    // ===================== TMA producer (converged warp, elected lane issues) =====================
    for (int g = 0; g < G; ++g) {
      const int hl = g / TQ, t = g % TQ;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      if (t == 0) {
        const int kvb = hl % C::kDqKVBufs, n = hl / C::kDqKVBufs;
        ptx::mbar_wait(&kv_empty[kvb], (n & 1) ^ 1);
        uint8_t* sK = sKV + kvb * 2 * C::kHeadBytes;
        uint8_t* sV = sK + C::kHeadBytes;
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&kv_full[kvb], 2 * C::kHeadBytes);
#pragma unroll
          for (int r = 0; r < C::kNKT; r += NKC) {
            ptx::tma_load_3d(sK + r * 128, &tmap_kv, &kv_full[kvb], D + h * 64, r, b);
            ptx::tma_load_3d(sV + r * 128, &tmap_kv, &kv_full[kvb], 2 * D + h * 64, r, b);
          }
        }
        __syncwarp();
      }
      const int qb = g & 1;
      ptx::mbar_wait(&t_empty[qb], ((g >> 1) & 1) ^ 1);
      uint8_t* st = sT + qb * 3 * C::kTileBytes;
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(&t_full[qb], 3 * C::kTileBytes);
        ptx::tma_load_3d(st, &tmap_q, &t_full[qb], h * 64, t * 128, b);
        ptx::tma_load_3d(st + C::kTileBytes, &tmap_do, &t_full[qb], h * 64, t * 128, b);
        ptx::tma_load_3d(st + 2 * C::kTileBytes, &tmap_o, &t_full[qb], h * 64, t * 128, b);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (converged warp; tcgen05.mma / commit under elect.sync, see attention2_sm100.cuh) ==========
    {
      constexpr uint32_t idesc_s = ptx::make_idesc_bf16(128, NKC);
      constexpr uint32_t idesc_o = ptx::make_idesc_bf16_ex(128, 64, 0, 1);
      constexpr uint32_t HI = ptx::kDescHiSw128;
      const uint32_t sT_u32 = ptx::smem_u32(sT), sKV_u32 = ptx::smem_u32(sKV);
      uint32_t n = 0;
      for (int g = 0; g < G; ++g) {
        const int hl = g / TQ, t = g % TQ;
        const int kvb = hl % C::kDqKVBufs, qb = g & 1;
        if (t == 0) ptx::mbar_wait(&kv_full[kvb], (hl / C::kDqKVBufs) & 1);
        ptx::mbar_wait(&t_full[qb], (g >> 1) & 1);
        ptx::tcgen05_fence_after();
        const uint32_t sq = sT_u32 + qb * 3 * C::kTileBytes;
        const uint32_t sdo = sq + C::kTileBytes;
        const uint32_t sk0 = sKV_u32 + kvb * 2 * C::kHeadBytes;
        const uint32_t sv0 = sk0 + C::kHeadBytes;
#pragma unroll 1
        for (int kc = 0; kc < NCH; ++kc) {
          const uint32_t sk = sk0 + kc * NKC * 128, sv = sv0 + kc * NKC * 128;
          if (ptx::elect_one()) {
            const uint32_t q_lo = ptx::desc_lo(sq), k_lo = ptx::desc_lo(sk), do_lo = ptx::desc_lo(sdo), v_lo = ptx::desc_lo(sv);
#pragma unroll
            for (int k = 0; k < 4; ++k) ptx::umma_bf16_ss_lohi(tmem_base, q_lo + 2 * k, k_lo + 2 * k, HI, idesc_s, k != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) ptx::umma_bf16_ss_lohi(tmem_base + NKC, do_lo + 2 * k, v_lo + 2 * k, HI, idesc_s, k != 0);
            ptx::umma_commit(sdp_ready);
          }
          __syncwarp();
          ptx::mbar_wait(ds_ready, n & 1);
          ++n;
          ptx::tcgen05_fence_after();
          if (kc == 0) {
            ptx::mbar_wait(dq_free, (g & 1) ^ 1);
            ptx::tcgen05_fence_after();
          }
          if (ptx::elect_one()) {
            const uint32_t k_lo = ptx::desc_lo(sk);
#pragma unroll
            for (int j = 0; j < NKC / 16; ++j)
              ptx::umma_bf16_ts_lohi(tmem_base + 2 * NKC, tmem_base + bwd_pcol<NKC, NCH>(j), k_lo + j * (2048 >> 4), HI, idesc_o,
                                     (kc | j) != 0 ? 1u : 0u);
          }
          __syncwarp();
        }
        if (ptx::elect_one()) {
          ptx::umma_commit(dq_ready);
          ptx::umma_commit(&t_empty[qb]);
          if (t == TQ - 1) ptx::umma_commit(&kv_empty[kvb]);
        }
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===================== compute warps =====================
    const int cw = warp - 4;
    const int quarter = warp & 3;
    const int wg = cw >> 2;
    const int rit = quarter * 32 + lane;
    const float c = p.scale * 1.4426950408889634f;
    const bool issuer = (threadIdx.x == 128);
    uint32_t n = 0;
    for (int g = 0; g < G; ++g) {
      const int hl = g / TQ, t = g % TQ;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      const int qb = g & 1;
      const int row = t * 128 + rit;
      const bool warp_active = (t * 128 + quarter * 32) < L;
      const uint32_t tb = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
      // ---- delta = rowsum(dO o O) from the TMA-staged tiles, lse from global
      ptx::mbar_wait(&t_full[qb], (g >> 1) & 1);
      float delta = 0.f, lse = 0.f;
      if (warp_active) {
        const uint8_t* sdo = sT + qb * 3 * C::kTileBytes + C::kTileBytes;
        const uint8_t* so = sdo + C::kTileBytes;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t off = (uint32_t)(rit * 128 + ((j ^ (rit & 7)) << 4));
          const uint4 a = *reinterpret_cast<const uint4*>(sdo + off);
          const uint4 o = *reinterpret_cast<const uint4*>(so + off);
          const uint32_t au[4] = {a.x, a.y, a.z, a.w}, ou[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float2 x = ptx::unpack_bf16x2(au[q]), y = ptx::unpack_bf16x2(ou[q]);
            delta = fmaf(x.x, y.x, delta);
            delta = fmaf(x.y, y.y, delta);
          }
        }
        if (row < L) {
          lse = p.lse2[(long long)bh * L + row];
          if (wg == 0) p.delta_out[(long long)bh * L + row] = delta;
        }
      }
#pragma unroll 1
      for (int kc = 0; kc < NCH; ++kc) {
        ptx::mbar_wait(sdp_ready, n & 1);
        ++n;
        ptx::tcgen05_fence_after();
        if (warp_active) {
          const int kbase = kc * NKC;      // global key index of chunk column 0
          if (wg == 0) {
#pragma unroll 1
            for (int c0 = 0; c0 + 32 <= C::kCA; c0 += 32)
              dq_chunk<32>(tb + c0, tb + NKC + c0, tb + (c0 >> 1), L - kbase - c0, c, lse, delta);
            if constexpr (C::kCA % 32 != 0) {
              constexpr int c0 = (C::kCA / 32) * 32;
              dq_chunk<16>(tb + c0, tb + NKC + c0, tb + (c0 >> 1), L - kbase - c0, c, lse, delta);
            }
          } else {
#pragma unroll 1
            for (int c0 = NKC - 32; c0 >= C::kCA; c0 -= 32)
              dq_chunk<32>(tb + c0, tb + NKC + c0, tb + (NKC - ((NKC - c0) >> 1)), L - kbase - c0, c, lse, delta);
          }
          ptx::tmem_st_wait();
        }
        ptx::tcgen05_fence_before();
        ptx::mbar_arrive(ds_ready);
      }
      // ---- epilogue: dQ * scale -> bf16 -> staging -> TMA store into the q block of dqkv
      ptx::mbar_wait(dq_ready, g & 1);
      ptx::tcgen05_fence_after();
      uint32_t r[32];
      if (warp_active) {
        tm_ld32(tb + 2 * NKC + wg * 32, r);
        ptx::tmem_ld_wait();
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(dq_free);
      if (issuer) ptx::bulk_wait_read<0>();
      ptx::named_bar_sync(1, 256);
      if (warp_active) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 u;
          u.x = ptx::pack_bf16x2(__uint_as_float(r[8 * j]) * p.scale, __uint_as_float(r[8 * j + 1]) * p.scale);
          u.y = ptx::pack_bf16x2(__uint_as_float(r[8 * j + 2]) * p.scale, __uint_as_float(r[8 * j + 3]) * p.scale);
          u.z = ptx::pack_bf16x2(__uint_as_float(r[8 * j + 4]) * p.scale, __uint_as_float(r[8 * j + 5]) * p.scale);
          u.w = ptx::pack_bf16x2(__uint_as_float(r[8 * j + 6]) * p.scale, __uint_as_float(r[8 * j + 7]) * p.scale);
          *reinterpret_cast<uint4*>(sOut + rit * 128 + (((wg * 4 + j) ^ (rit & 7)) << 4)) = u;
        }
      }
      ptx::fence_proxy_async();
      ptx::named_bar_sync(1, 256);
      if (issuer) {
        ptx::tma_store_3d(&tmap_out, sOut, h * 64, t * 128, b);
        ptx::bulk_commit();
      }
    }
    if (issuer) ptx::bulk_wait<0>();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  pdl_trigger();                                         // PDL: late on purpose (see gemm2_sm100.cuh)
  if (warp == 2) ptx::tmem_dealloc<C::kDqTmemCols>(tmem_base);
}

// =====================================================================================================================
// dK / dV kernel
// =====================================================================================================================
template <int NKC, int NCH>
__global__ void __launch_bounds__(kA2Threads, 1)
attn2_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmap_kt, const __grid_constant__ CUtensorMap tmap_q,
                     const __grid_constant__ CUtensorMap tmap_do, const __grid_constant__ CUtensorMap tmap_out,
                     const Attn2Params p) {
  using C = Attn2BwdCfg<NKC, NCH>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sQD = smem;                                          // [kDkvQDBufs][Q | dO]   (all query rows of the head)
  uint8_t* sT = sQD + 2 * C::kDkvQDBufs * C::kHeadBytes;        // [2 tiles][K_j | V_j]
  uint8_t* sOut = sT + 4 * C::kTileBytes;                       // [2] dK / dV staging
  float* sLse = reinterpret_cast<float*>(sOut + 2 * C::kTileBytes);
  float* sDelta = sLse + C::kNKT;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDelta + C::kNKT);
  uint64_t* qd_full = bars;            // [2]
  uint64_t* qd_empty = bars + 2;       // [2]
  uint64_t* t_full = bars + 4;         // [2]
  uint64_t* t_empty = bars + 6;        // [2]
  uint64_t* sdp_ready = bars + 8;
  uint64_t* pds_ready = bars + 9;
  uint64_t* out_ready = bars + 10;
  uint64_t* out_free = bars + 11;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int L = p.L, H = p.H;
  const int TK = (L + 127) / 128;
  const int total_heads = p.B * H;
  const int my_heads = (total_heads - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int G = my_heads * TK;
  const int D = H * 64;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_kt); ptx::prefetch_tmap(&tmap_q); ptx::prefetch_tmap(&tmap_do); ptx::prefetch_tmap(&tmap_out);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&qd_full[i], 1);
      ptx::mbar_init(&qd_empty[i], 1);
      ptx::mbar_init(&t_full[i], 1);
      ptx::mbar_init(&t_empty[i], 1);
    }
    ptx::mbar_init(sdp_ready, 1);
    ptx::mbar_init(pds_ready, 256);
    ptx::mbar_init(out_ready, 1);
    ptx::mbar_init(out_free, 256);
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc<C::kDkvTmemCols>(tmem_slot);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                            // PDL (common.cuh): no global-memory access above this line

  if (warp == 0) {
    // ===================== TMA producer (converged warp, elected lane issues) =====================
    for (int g = 0; g < G; ++g) {
      const int hl = g / TK, j = g % TK;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      if (j == 0) {
        const int qdb = hl % C::kDkvQDBufs, n = hl / C::kDkvQDBufs;
        ptx::mbar_wait(&qd_empty[qdb], (n & 1) ^ 1);
        uint8_t* sQ = sQD + qdb * 2 * C::kHeadBytes;
        uint8_t* sdO = sQ + C::kHeadBytes;
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&qd_full[qdb], 2 * C::kHeadBytes);
#pragma unroll
          for (int r = 0; r < C::kNKT; r += NKC) {
            ptx::tma_load_3d(sQ + r * 128, &tmap_q, &qd_full[qdb], h * 64, r, b);
            ptx::tma_load_3d(sdO + r * 128, &tmap_do, &qd_full[qdb], h * 64, r, b);
          }
        }
        __syncwarp();
      }
      const int kb = g & 1;
      ptx::mbar_wait(&t_empty[kb], ((g >> 1) & 1) ^ 1);
      uint8_t* st = sT + kb * 2 * C::kTileBytes;
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(&t_full[kb], 2 * C::kTileBytes);
        ptx::tma_load_3d(st, &tmap_kt, &t_full[kb], D + h * 64, j * 128, b);
        ptx::tma_load_3d(st + C::kTileBytes, &tmap_kt, &t_full[kb], 2 * D + h * 64, j * 128, b);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (converged warp; tcgen05.mma / commit under elect.sync) =====================
    {
      constexpr uint32_t idesc_s = ptx::make_idesc_bf16(128, NKC);
      constexpr uint32_t idesc_o = ptx::make_idesc_bf16_ex(128, 64, 0, 1);
      constexpr uint32_t HI = ptx::kDescHiSw128;
      const uint32_t sT_u32 = ptx::smem_u32(sT), sQD_u32 = ptx::smem_u32(sQD);
      uint32_t n = 0;
      for (int g = 0; g < G; ++g) {
        const int hl = g / TK, j = g % TK;
        const int qdb = hl % C::kDkvQDBufs, kb = g & 1;
        if (j == 0) ptx::mbar_wait(&qd_full[qdb], (hl / C::kDkvQDBufs) & 1);
        ptx::mbar_wait(&t_full[kb], (g >> 1) & 1);
        ptx::mbar_wait(out_free, (g & 1) ^ 1);          // the previous tile's dK / dV accumulators have been read out
        ptx::tcgen05_fence_after();
        const uint32_t skj = sT_u32 + kb * 2 * C::kTileBytes;
        const uint32_t svj = skj + C::kTileBytes;
        const uint32_t sq0 = sQD_u32 + qdb * 2 * C::kHeadBytes;
        const uint32_t sdo0 = sq0 + C::kHeadBytes;
#pragma unroll 1
        for (int qc = 0; qc < NCH; ++qc) {
          const uint32_t sq = sq0 + qc * NKC * 128, sdo = sdo0 + qc * NKC * 128;
          if (ptx::elect_one()) {
            const uint32_t k_lo = ptx::desc_lo(skj), q_lo = ptx::desc_lo(sq), v_lo = ptx::desc_lo(svj), do_lo = ptx::desc_lo(sdo);
#pragma unroll
            for (int k = 0; k < 4; ++k) ptx::umma_bf16_ss_lohi(tmem_base, k_lo + 2 * k, q_lo + 2 * k, HI, idesc_s, k != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) ptx::umma_bf16_ss_lohi(tmem_base + NKC, v_lo + 2 * k, do_lo + 2 * k, HI, idesc_s, k != 0);
            ptx::umma_commit(sdp_ready);
          }
          __syncwarp();
          ptx::mbar_wait(pds_ready, n & 1);
          ++n;
          ptx::tcgen05_fence_after();
          if (ptx::elect_one()) {
            const uint32_t q_lo = ptx::desc_lo(sq), do_lo = ptx::desc_lo(sdo);
#pragma unroll
            for (int jj = 0; jj < NKC / 16; ++jj) {
              const uint32_t pc = bwd_pcol<NKC, NCH>(jj);
              const uint32_t acc = (qc | jj) != 0 ? 1u : 0u;
              ptx::umma_bf16_ts_lohi(tmem_base + C::kDvCol, tmem_base + pc, do_lo + jj * (2048 >> 4), HI, idesc_o, acc);
              ptx::umma_bf16_ts_lohi(tmem_base + C::kDkCol, tmem_base + NKC + pc, q_lo + jj * (2048 >> 4), HI, idesc_o, acc);
            }
          }
          __syncwarp();
        }
        if (ptx::elect_one()) {
          ptx::umma_commit(out_ready);
          ptx::umma_commit(&t_empty[kb]);
          if (j == TK - 1) ptx::umma_commit(&qd_empty[qdb]);
        }
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===================== compute warps =====================
    const int cw = warp - 4;
    const int quarter = warp & 3;
    const int wg = cw >> 2;
    const int rit = quarter * 32 + lane;
    const int ctid = threadIdx.x - 128;
    const float c = p.scale * 1.4426950408889634f;
    const bool issuer = ((cw & 3) == 0) && lane == 0;
    uint8_t* stage = sOut + wg * C::kTileBytes;
    uint32_t n = 0;
    for (int g = 0; g < G; ++g) {
      const int hl = g / TK, j = g % TK;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      const bool warp_active = (j * 128 + quarter * 32) < L;
      const uint32_t tb = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
      if (j == 0) {
        // per-head lse / delta of every query (columns of S^T) -> smem
        ptx::named_bar_sync(1, 256);
        for (int i = ctid; i < C::kNKT; i += 256) {
          sLse[i] = i < L ? p.lse2[(long long)bh * L + i] : 0.f;
          sDelta[i] = i < L ? p.delta_in[(long long)bh * L + i] : 0.f;
        }
        ptx::named_bar_sync(1, 256);
      }
#pragma unroll 1
      for (int qc = 0; qc < NCH; ++qc) {
        ptx::mbar_wait(sdp_ready, n & 1);
        ++n;
        ptx::tcgen05_fence_after();
        if (warp_active) {
          const int qbase = qc * NKC;
          if (wg == 0) {
#pragma unroll 1
            for (int c0 = 0; c0 + 32 <= C::kCA; c0 += 32)
              dkv_chunk<32>(tb + c0, tb + NKC + c0, tb + (c0 >> 1), tb + NKC + (c0 >> 1), L - qbase - c0, c, sLse + qbase + c0,
                            sDelta + qbase + c0);
            if constexpr (C::kCA % 32 != 0) {
              constexpr int c0 = (C::kCA / 32) * 32;
              dkv_chunk<16>(tb + c0, tb + NKC + c0, tb + (c0 >> 1), tb + NKC + (c0 >> 1), L - qbase - c0, c, sLse + qbase + c0,
                            sDelta + qbase + c0);
            }
          } else {
#pragma unroll 1
            for (int c0 = NKC - 32; c0 >= C::kCA; c0 -= 32) {
              const uint32_t pc = (uint32_t)(NKC - ((NKC - c0) >> 1));
              dkv_chunk<32>(tb + c0, tb + NKC + c0, tb + pc, tb + NKC + pc, L - qbase - c0, c, sLse + qbase + c0, sDelta + qbase + c0);
            }
          }
          ptx::tmem_st_wait();
        }
        ptx::tcgen05_fence_before();
        ptx::mbar_arrive(pds_ready);
      }
      // ---- epilogue: warpgroup A stores dK_j * scale, warpgroup B stores dV_j
      ptx::mbar_wait(out_ready, g & 1);
      ptx::tcgen05_fence_after();
      uint32_t r[64];
      if (warp_active) {
        const uint32_t col = wg == 0 ? C::kDkCol : C::kDvCol;
        tm_ld32(tb + col, r);
        tm_ld32(tb + col + 32, r + 32);
        ptx::tmem_ld_wait();
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(out_free);
      if (issuer) ptx::bulk_wait_read<0>();
      ptx::named_bar_sync(2 + wg, 128);
      if (warp_active) {
        const float f = wg == 0 ? p.scale : 1.0f;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          uint4 u;
          u.x = ptx::pack_bf16x2(__uint_as_float(r[8 * q]) * f, __uint_as_float(r[8 * q + 1]) * f);
          u.y = ptx::pack_bf16x2(__uint_as_float(r[8 * q + 2]) * f, __uint_as_float(r[8 * q + 3]) * f);
          u.z = ptx::pack_bf16x2(__uint_as_float(r[8 * q + 4]) * f, __uint_as_float(r[8 * q + 5]) * f);
          u.w = ptx::pack_bf16x2(__uint_as_float(r[8 * q + 6]) * f, __uint_as_float(r[8 * q + 7]) * f);
          *reinterpret_cast<uint4*>(stage + rit * 128 + ((q ^ (rit & 7)) << 4)) = u;
        }
      }
      ptx::fence_proxy_async();
      ptx::named_bar_sync(2 + wg, 128);
      if (issuer) {
        ptx::tma_store_3d(&tmap_out, stage, (wg == 0 ? D : 2 * D) + h * 64, j * 128, b);
        ptx::bulk_commit();
      }
    }
    if (issuer) ptx::bulk_wait<0>();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  pdl_trigger();                                         // PDL: late on purpose (see gemm2_sm100.cuh)
  if (warp == 2) ptx::tmem_dealloc<C::kDkvTmemCols>(tmem_base);
}

}  // namespace peft
