// tcgen05 / TMEM / TMA multi-head attention backward, SOFTWARE-PIPELINED (L <= 272, head_dim 64).
//
// Same two kernels and the same arithmetic as attention2_bwd_sm100.cuh (dQ kernel with TMEM lane = query row, dK/dV kernel
// with TMEM lane = key row; S and dP as SS products, bf16 P / dS written back in place, TS products for the outputs), but the
// score columns of a tile are walked in 64-wide ITEMS that rotate through THREE TMEM buffers (S | dP = 128 columns each),
// and the two compute warpgroups take alternate items.  The MMA warp issues the SS products of item i + 2 as soon as the TS
// products of item i - 1 are in the queue, so the tensor core computes S / dP of the items ahead and dQ (dK, dV) of the items
// behind while a warpgroup exponentiates the current one — in the attention2 kernels the tensor core and the MUFU / FMA pipes
// took turns (measured there: ~6 k clk per 128-row tile against ~1.2 k clk of MMA work and ~1.7 k clk of MUFU work).
//   TMEM (512 columns):  [0,384) three S|dP buffers,  [384,512) dQ accumulators of two consecutive tiles (dQ kernel)
//                                                      or dV | dK of the current key tile (dK/dV kernel)
// Results leave straight from registers (one thread = one row = 128 contiguous bytes), so no staging tile, no proxy fence and
// no CTA-wide barrier sits between the warpgroups.
//
// replaces: the autograd backward of FS/kadaptation.py:223-235 / FS/model.py:990-1002.
#pragma once
#include "attention2_bwd_sm100.cuh"

namespace peft {

struct Attn3Params {
  const float* lse2;        // [B*H, L]  log2-domain log-sum-exp of the forward
  const float* delta_in;    // dK/dV kernel: [B*H, L]
  float* delta_out;         // dQ kernel:    [B*H, L]
  __nv_bfloat16* dqkv;      // [B*L, lddqkv]: dq | dk | dv column blocks of width D = 64 H
  long long lddqkv;
  int B, H, L;
  float scale;
  int prefetch;                // L2 prefetch of operand boxes ahead of their loads (env PEFT_ATTN_PREFETCH=1; off by default: measured no gain)
  int flags;                   // bit 0: 32-column math steps without load pipelining (A/B switch, env PEFT_ATTN3_FLAGS)
  unsigned long long* stats;   // diagnostics (PEFT_ATTN_STATS=1): per CTA 16 counters, see tests/bench_ops.py attn_bwd_stats
};

template <int NKT>   // padded number of keys (= queries) of a head: 64, 208 or 272
struct Attn3Cfg {
  static_assert(NKT % 16 == 0 && NKT >= 64 && NKT <= 272, "padded sequence length");
  static constexpr int kNCH = (NKT + 63) / 64;                 // items per tile
  static constexpr int kWLast = NKT - 64 * (kNCH - 1);         // width of the last item (64 or 16)
  static_assert(kWLast == 64 || kWLast == 16, "last item width");
  static constexpr int kNBuf = 3;
  static constexpr int kAccCol = kNBuf * 128;
  static constexpr int kBox = NKT <= 256 ? NKT : NKT / 2;      // rows per TMA box of a per-head operand
  static constexpr int kHeadBytes = NKT * 128;
  static constexpr int kTileBytes = 128 * 128;
  // dQ kernel: [bufs][K | V] | [2][Q | dO | O] | barriers
  static constexpr int kDqFixed = 6 * kTileBytes + 512 + 1024;
  static constexpr int kDqKVBufs = (4 * kHeadBytes + kDqFixed <= 232448) ? 2 : 1;
  static constexpr int kDqSmemBytes = 2 * kDqKVBufs * kHeadBytes + kDqFixed;
  // dK/dV kernel: [2][Q | dO] | [2][K_j | V_j] | [2 warpgroups][lse | delta] | barriers
  static constexpr int kDkvFixed = 4 * kTileBytes + 4 * NKT * 4 + 512 + 1024;
  static_assert(4 * kHeadBytes + kDkvFixed <= 232448, "dK/dV kernel: two heads of Q / dO must fit");
  static constexpr int kDkvSmemBytes = 4 * kHeadBytes + kDkvFixed;
};

// One ITEM (W = 64 or 16 score columns) of the dQ kernel in 16-column steps with the TMEM loads of step k + 1 in flight while
// step k is exponentiated (tcgen05.wait::ld waits for every outstanding load, so the next loads are issued right after it).
// dS (bf16) overwrites the S columns in place: step k writes packed columns [8k, 8k + 8), all below anything still unread.
template <int W>
__device__ __forceinline__ void a3_dq_item(uint32_t tS, int valid, float c, float lse, float delta) {
  constexpr int NC = W / 16;
  uint32_t s[2][16], d[2][16];
  tm_ld16(tS, s[0]);
  tm_ld16(tS + 64, d[0]);
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    ptx::tmem_ld_wait();
    if (k + 1 < NC) {
      tm_ld16(tS + 16 * (k + 1), s[(k + 1) & 1]);
      tm_ld16(tS + 64 + 16 * (k + 1), d[(k + 1) & 1]);
    }
    uint32_t pk[8];
    const int v = valid - 16 * k;
#pragma unroll
    for (int j = 0; j < 16; j += 2) {
      const float p0 = (j < v) ? ptx::ex2_approx(fmaf(__uint_as_float(s[k & 1][j]), c, -lse)) : 0.f;
      const float p1 = (j + 1 < v) ? ptx::ex2_approx(fmaf(__uint_as_float(s[k & 1][j + 1]), c, -lse)) : 0.f;
      pk[j >> 1] = ptx::pack_bf16x2(p0 * (__uint_as_float(d[k & 1][j]) - delta), p1 * (__uint_as_float(d[k & 1][j + 1]) - delta));
    }
    tm_st8(tS + 8 * k, pk);
  }
}

// Same for the dK/dV kernel (columns = queries, per-column lse / delta from shared memory): P^T goes to the S^T columns, dS^T to
// the dP^T columns.
template <int W>
__device__ __forceinline__ void a3_dkv_item(uint32_t tS, int valid, float c, const float* __restrict__ lse,
                                            const float* __restrict__ delta) {
  constexpr int NC = W / 16;
  uint32_t s[2][16], d[2][16];
  tm_ld16(tS, s[0]);
  tm_ld16(tS + 64, d[0]);
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    ptx::tmem_ld_wait();
    if (k + 1 < NC) {
      tm_ld16(tS + 16 * (k + 1), s[(k + 1) & 1]);
      tm_ld16(tS + 64 + 16 * (k + 1), d[(k + 1) & 1]);
    }
    uint32_t pp[8], pd[8];
    const int v = valid - 16 * k;
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      const float4 l4 = *reinterpret_cast<const float4*>(lse + 16 * k + j);
      const float4 d4 = *reinterpret_cast<const float4*>(delta + 16 * k + j);
      const float lv[4] = {l4.x, l4.y, l4.z, l4.w}, dv[4] = {d4.x, d4.y, d4.z, d4.w};
      float pv[4], sv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        pv[q] = (j + q < v) ? ptx::ex2_approx(fmaf(__uint_as_float(s[k & 1][j + q]), c, -lv[q])) : 0.f;
        sv[q] = pv[q] * (__uint_as_float(d[k & 1][j + q]) - dv[q]);
      }
      pp[j >> 1] = ptx::pack_bf16x2(pv[0], pv[1]);
      pp[(j >> 1) + 1] = ptx::pack_bf16x2(pv[2], pv[3]);
      pd[j >> 1] = ptx::pack_bf16x2(sv[0], sv[1]);
      pd[(j >> 1) + 1] = ptx::pack_bf16x2(sv[2], sv[3]);
    }
    tm_st8(tS + 8 * k, pp);
    tm_st8(tS + 64 + 8 * k, pd);
  }
}

// One chunk of W query columns of the dK/dV kernel, register-lean form of dkv_chunk (attention2_bwd_sm100.cuh): P^T replaces
// S^T in its registers and is packed + stored before dS^T is formed in the dP^T registers (peak 2 W + W / 2 live values).
template <int W>
__device__ __forceinline__ void a3_dkv_chunk(uint32_t ts, uint32_t tdp, uint32_t tout_p, uint32_t tout_ds, int valid, float c,
                                             const float* __restrict__ lse, const float* __restrict__ delta) {
  uint32_t s[W], d[W], pk[W / 2];
  tm_ld<W>(ts, s);
  tm_ld<W>(tdp, d);
  ptx::tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < W; j += 4) {
    const float4 l4 = *reinterpret_cast<const float4*>(lse + j);
    const float lv[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
    for (int q = 0; q < 4; ++q)
      s[j + q] = __float_as_uint((j + q < valid) ? ptx::ex2_approx(fmaf(__uint_as_float(s[j + q]), c, -lv[q])) : 0.f);
  }
#pragma unroll
  for (int j = 0; j < W; j += 2) pk[j >> 1] = ptx::pack_bf16x2(__uint_as_float(s[j]), __uint_as_float(s[j + 1]));
  tm_st_packed<W>(tout_p, pk);
#pragma unroll
  for (int j = 0; j < W; j += 4) {
    const float4 d4 = *reinterpret_cast<const float4*>(delta + j);
    const float dv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) d[j + q] = __float_as_uint(__uint_as_float(s[j + q]) * (__uint_as_float(d[j + q]) - dv[q]));
  }
#pragma unroll
  for (int j = 0; j < W; j += 2) pk[j >> 1] = ptx::pack_bf16x2(__uint_as_float(d[j]), __uint_as_float(d[j + 1]));
  tm_st_packed<W>(tout_ds, pk);
}

// =====================================================================================================================
// dQ kernel
// =====================================================================================================================
template <int NKT>
__global__ void __launch_bounds__(kA2Threads, 1)
attn3_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                    const __grid_constant__ CUtensorMap tmap_do, const __grid_constant__ CUtensorMap tmap_o, const Attn3Params p) {
  using C = Attn3Cfg<NKT>;
  constexpr int NCH = C::kNCH, NB = C::kNBuf;
#if PEFT_GEMM_DIAG        // cycle counters and the diagnostic A/B flags only in the diagnostics build (make DIAG=1)
  unsigned long long* const dstats = p.stats;
  const int dflags = p.flags;
#else
  constexpr unsigned long long* dstats = nullptr;
  constexpr int dflags = 0;
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sKV = smem;                                          // [kDqKVBufs][K | V]
  uint8_t* sT = sKV + 2 * C::kDqKVBufs * C::kHeadBytes;         // [2 tiles][Q | dO | O]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sT + 6 * C::kTileBytes);
  uint64_t* kv_full = bars;            // [2]
  uint64_t* kv_empty = bars + 2;       // [2]
  uint64_t* t_full = bars + 4;         // [2]
  uint64_t* t_empty = bars + 6;        // [2]
  uint64_t* sdp_ready = bars + 8;      // [3]  S / dP of the item in buffer b are complete
  uint64_t* ds_ready = bars + 11;      // [3]  bf16 dS of the item in buffer b is in place
  uint64_t* dq_ready = bars + 14;      // [2]  dQ accumulator a holds the whole tile
  uint64_t* dq_free = bars + 16;       // [2]  ... and has been read out
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int L = p.L, H = p.H;
  const int TQ = (L + 127) / 128;
  const int total_heads = p.B * H;
  const int my_heads = (total_heads - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int G = my_heads * TQ;
  const int N = G * NCH;               // items of this CTA
  const int D = H * 64;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_q); ptx::prefetch_tmap(&tmap_kv); ptx::prefetch_tmap(&tmap_do); ptx::prefetch_tmap(&tmap_o);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&kv_full[i], 1);
      ptx::mbar_init(&kv_empty[i], 1);
      ptx::mbar_init(&t_full[i], 1);
      ptx::mbar_init(&t_empty[i], 1);
      ptx::mbar_init(&dq_ready[i], 1);
      ptx::mbar_init(&dq_free[i], 128);
    }
    for (int i = 0; i < NB; ++i) {
      ptx::mbar_init(&sdp_ready[i], 1);
      ptx::mbar_init(&ds_ready[i], 128);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc<512>(tmem_slot);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                            // PDL (common.cuh): no global-memory access above this line

  if (warp == 0) {
    // ===================== TMA producer (converged warp, elected lane issues) =====================
    // L2 prefetch of the operands of tile g2 (attention2_sm100.cuh: tma_prefetch_3d), four tiles ahead of their loads
    auto prefetch = [&](int g2) {
      if (g2 >= G || !p.prefetch) return;
      const int hl2 = g2 / TQ, t2 = g2 % TQ;
      const int bh2 = (int)blockIdx.x + hl2 * (int)gridDim.x;
      const int b2 = bh2 / H, h2 = bh2 % H;
      if (ptx::elect_one()) {
        if (t2 == 0) {
#pragma unroll
          for (int r = 0; r < NKT; r += C::kBox) {
            ptx::tma_prefetch_3d(&tmap_kv, D + h2 * 64, r, b2);
            ptx::tma_prefetch_3d(&tmap_kv, 2 * D + h2 * 64, r, b2);
          }
        }
        ptx::tma_prefetch_3d(&tmap_q, h2 * 64, t2 * 128, b2);
        ptx::tma_prefetch_3d(&tmap_do, h2 * 64, t2 * 128, b2);
        ptx::tma_prefetch_3d(&tmap_o, h2 * 64, t2 * 128, b2);
      }
      __syncwarp();
    };
    prefetch(2);
    prefetch(3);
    for (int g = 0; g < G; ++g) {
      const int hl = g / TQ, t = g % TQ;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      prefetch(g + 4);
      if (t == 0) {
        const int kvb = hl % C::kDqKVBufs, n = hl / C::kDqKVBufs;
        ptx::mbar_wait(&kv_empty[kvb], (n & 1) ^ 1);
        uint8_t* sK = sKV + kvb * 2 * C::kHeadBytes;
        uint8_t* sV = sK + C::kHeadBytes;
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&kv_full[kvb], 2 * C::kHeadBytes);
#pragma unroll
          for (int r = 0; r < NKT; r += C::kBox) {
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
    // ===================== MMA issuer (converged warp; tcgen05.mma / commit under elect.sync) =====================
    constexpr uint32_t idesc_s64 = ptx::make_idesc_bf16(128, 64);
    constexpr uint32_t idesc_sl = ptx::make_idesc_bf16(128, C::kWLast);
    constexpr uint32_t idesc_o = ptx::make_idesc_bf16_ex(128, 64, 0, 1);
    constexpr uint32_t HI = ptx::kDescHiSw128;
    const uint32_t sT_u32 = ptx::smem_u32(sT), sKV_u32 = ptx::smem_u32(sKV);
    const int head_items = TQ * NCH;
    // S = Q_t K_c^T and dP = dO_t V_c^T of item k into buffer k % 3
    auto issue_ss = [&](int k) {
      const int g = k / NCH, kc = k % NCH;
      const int hl = g / TQ, t = g % TQ;
      const int kvb = hl % C::kDqKVBufs, qb = g & 1;
      if (kc == 0) {
        if (t == 0) ptx::mbar_wait(&kv_full[kvb], (hl / C::kDqKVBufs) & 1);
        ptx::mbar_wait(&t_full[qb], (g >> 1) & 1);
        ptx::tcgen05_fence_after();
      }
      const uint32_t sq = sT_u32 + qb * 3 * C::kTileBytes;
      const uint32_t sdo = sq + C::kTileBytes;
      const uint32_t sk = sKV_u32 + kvb * 2 * C::kHeadBytes + kc * 64 * 128;
      const uint32_t sv = sk + C::kHeadBytes;
      const uint32_t tS = tmem_base + (uint32_t)((k % NB) * 128);
      const uint32_t idesc = (kc == NCH - 1) ? idesc_sl : idesc_s64;
      if (ptx::elect_one()) {
        const uint32_t q_lo = ptx::desc_lo(sq), k_lo = ptx::desc_lo(sk), do_lo = ptx::desc_lo(sdo), v_lo = ptx::desc_lo(sv);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) ptx::umma_bf16_ss_lohi(tS, q_lo + 2 * kk, k_lo + 2 * kk, HI, idesc, kk != 0);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) ptx::umma_bf16_ss_lohi(tS + 64, do_lo + 2 * kk, v_lo + 2 * kk, HI, idesc, kk != 0);
        ptx::umma_commit(&sdp_ready[k % NB]);
      }
      __syncwarp();
    };
    // Item k may be issued once (a) the TS products of item k - 3 (its TMEM buffer) are in the queue, (b) the tile slot of its
    // Q / dO is free, i.e. the tile two before has been finished, and (c) with a single K / V buffer the previous head has been
    // finished — every blocking wait inside issue_ss() is then satisfied without any further action of this warp.
    // The Q / dO / O slot of a tile is handed back to the TMA producer EARLY: once the tile's last SS product has been issued
    // (it is, with the TS products of tile-local item NCH - 4) and both warpgroups have taken delta from the dO / O tiles (they
    // have, once items 0 and 1 of the tile have come back) — i.e. together with the TS products of tile-local item kRel, not
    // at the end of the tile.  The load of tile g + 2 then has more than a whole tile time to land; released at the end of tile g
    // it was the critical path (r02 counters: the MMA warp spent 30 % of its time waiting for operands inside issue_ss).
    constexpr int kRel = (NCH >= 4) ? NCH - 3 : (NCH >= 2 ? 1 : 0);
    int next_ss = 0;
    long long w_ss = 0, w_ds = 0, w_free = 0, tc0 = 0;
    const long long tbeg = clock64();
    auto advance = [&](int done) {
      while (next_ss < N && next_ss < done + NB && (next_ss / NCH - 2) * NCH + kRel + 1 <= done &&
             (next_ss / head_items - (C::kDqKVBufs - 1)) * head_items <= done) {      // K / V slot: head hl - kDqKVBufs is finished
        if (dstats) tc0 = clock64();
        issue_ss(next_ss);
        if (dstats) w_ss += clock64() - tc0;
        ++next_ss;
      }
    };
    advance(0);
    for (int i = 0; i < N; ++i) {
      const int g = i / NCH, kc = i % NCH;
      const int hl = g / TQ, t = g % TQ;
      const int kvb = hl % C::kDqKVBufs, qb = g & 1, a = g & 1;
      if (dstats) tc0 = clock64();
      ptx::mbar_wait(&ds_ready[i % NB], (i / NB) & 1);
      if (dstats) { const long long n_ = clock64(); w_ds += n_ - tc0; tc0 = n_; }
      if (kc == 0) ptx::mbar_wait(&dq_free[a], ((g >> 1) & 1) ^ 1);
      if (dstats) w_free += clock64() - tc0;
      ptx::tcgen05_fence_after();
      const uint32_t sk = sKV_u32 + kvb * 2 * C::kHeadBytes + kc * 64 * 128;
      const uint32_t tS = tmem_base + (uint32_t)((i % NB) * 128);
      const uint32_t tD = tmem_base + (uint32_t)(C::kAccCol + 64 * a);
      // (both trip counts are compile-time constants: a rolled loop around tcgen05.mma costs ~210 clk per instruction — the
      // descriptors then reach the uniform registers through an R2UR waterfall, bench_ops.py mma_probe)
      if (ptx::elect_one()) {
        const uint32_t k_lo = ptx::desc_lo(sk);
        if (C::kWLast == 64 || kc != NCH - 1) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            ptx::umma_bf16_ts_lohi(tD, tS + 8 * j, k_lo + j * (2048 >> 4), HI, idesc_o, (kc | j) != 0 ? 1u : 0u);
        } else {
#pragma unroll
          for (int j = 0; j < C::kWLast / 16; ++j)
            ptx::umma_bf16_ts_lohi(tD, tS + 8 * j, k_lo + j * (2048 >> 4), HI, idesc_o, 1u);
        }
        if (kc == kRel) ptx::umma_commit(&t_empty[qb]);     // the tile's last SS products went out with item kRel - 1 at the latest
        if (kc == NCH - 1) {
          ptx::umma_commit(&dq_ready[a]);
          if (t == TQ - 1) ptx::umma_commit(&kv_empty[kvb]);
        }
      }
      __syncwarp();
      advance(i + 1);
    }
    if (dstats && lane == 0) {
      unsigned long long* o = dstats + blockIdx.x * 16;
      o[0] = clock64() - tbeg; o[1] = w_ss; o[2] = w_ds; o[3] = w_free;
    }
  } else if (warp >= 4) {
    // ===================== compute warps: warpgroup w takes the items i = w, w + 2, ... =====================
    const int quarter = warp & 3;
    const int wg = (warp - 4) >> 2;
    const int rit = quarter * 32 + lane;
    const float c = p.scale * 1.4426950408889634f;
    const uint32_t tb = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    long long c_delta = 0, c_wait = 0, c_math = 0, c_epi = 0, tc0 = clock64();
    const long long tbeg = tc0;
    auto lap = [&](long long& acc) { if (dstats) { const long long n_ = clock64(); acc += n_ - tc0; tc0 = n_; } };
    int cur_g = -1;
    float delta = 0.f, lse = 0.f;
    for (int i = wg; i < N; i += 2) {
      const int g = i / NCH, kc = i % NCH;
      const int hl = g / TQ, t = g % TQ;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int qb = g & 1;
      const int row = t * 128 + rit;
      const bool warp_active = (t * 128 + quarter * 32) < L;
      if (g != cur_g) {
        // ---- delta = rowsum(dO o O) from the TMA-staged tiles, lse from global
        cur_g = g;
        ptx::mbar_wait(&t_full[qb], (g >> 1) & 1);
        delta = 0.f; lse = 0.f;
        if (warp_active && !(dflags & 4)) {   // diagnostic bit 2: no delta (results wrong)
          const uint8_t* sdo = sT + qb * 3 * C::kTileBytes + C::kTileBytes;
          const uint8_t* so = sdo + C::kTileBytes;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t off = (uint32_t)(rit * 128 + ((j ^ (rit & 7)) << 4));
            const uint4 av = *reinterpret_cast<const uint4*>(sdo + off);
            const uint4 ov = *reinterpret_cast<const uint4*>(so + off);
            const uint32_t au[4] = {av.x, av.y, av.z, av.w}, ou[4] = {ov.x, ov.y, ov.z, ov.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float2 x = ptx::unpack_bf16x2(au[q]), y = ptx::unpack_bf16x2(ou[q]);
              delta = fmaf(x.x, y.x, delta);
              delta = fmaf(x.y, y.y, delta);
            }
          }
          if (row < L) lse = p.lse2[(long long)bh * L + row];
        }
      }
      if (kc == 0 && warp_active && row < L) p.delta_out[(long long)bh * L + row] = delta;
      const int buf = i % NB;
      lap(c_delta);
      ptx::mbar_wait(&sdp_ready[buf], (i / NB) & 1);
      ptx::tcgen05_fence_after();
      lap(c_wait);
      if (warp_active && !(dflags & 2)) {   // diagnostic bit 1: no math (results wrong)
        const uint32_t tS = tb + (uint32_t)(buf * 128);
        const int valid = L - kc * 64;
        if (C::kWLast == 64 || kc != NCH - 1) {
          if (dflags & 1) {
            dq_chunk<32>(tS, tS + 64, tS, valid, c, lse, delta);
            dq_chunk<32>(tS + 32, tS + 96, tS + 16, valid - 32, c, lse, delta);
          } else {
            a3_dq_item<64>(tS, valid, c, lse, delta);
          }
        } else {
          dq_chunk<16>(tS, tS + 64, tS, valid, c, lse, delta);
        }
        ptx::tmem_st_wait();
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(&ds_ready[buf]);
      lap(c_math);
      if (kc == NCH - 1) {
        // ---- epilogue of the tile (by the warpgroup that took its last item): dQ * scale -> bf16 -> global
        const int a = g & 1;
        const int b = bh / H, h = bh % H;
        ptx::mbar_wait(&dq_ready[a], (g >> 1) & 1);
        ptx::tcgen05_fence_after();
        __nv_bfloat16* dst = p.dqkv + ((long long)b * L + row) * p.lddqkv + h * 64;
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t r[32], pk[16];
          if (warp_active) {
            tm_ld32(tb + C::kAccCol + 64 * a + 32 * hf, r);
            ptx::tmem_ld_wait();
          }
          if (hf == 1) {
            ptx::tcgen05_fence_before();
            ptx::mbar_arrive(&dq_free[a]);
          }
          if (warp_active && row < L) {
#pragma unroll
            for (int q = 0; q < 16; ++q)
              pk[q] = ptx::pack_bf16x2(__uint_as_float(r[2 * q]) * p.scale, __uint_as_float(r[2 * q + 1]) * p.scale);
            uint4* d4 = reinterpret_cast<uint4*>(dst + 32 * hf);
#pragma unroll
            for (int q = 0; q < 4; ++q) d4[q] = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
          }
        }
        lap(c_epi);
      }
    }
    if (dstats && (warp & 3) == 0 && lane == 0) {
      unsigned long long* o = dstats + blockIdx.x * 16 + 4 + 6 * wg;
      o[0] = clock64() - tbeg; o[1] = c_delta; o[2] = c_wait; o[3] = c_math; o[4] = c_epi;
    }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  pdl_trigger();                                         // PDL: late on purpose (see gemm2_sm100.cuh)
  if (warp == 2) ptx::tmem_dealloc<512>(tmem_base);
}

// =====================================================================================================================
// dK / dV kernel
// =====================================================================================================================
template <int NKT>
__global__ void __launch_bounds__(kA2Threads, 1)
attn3_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmap_kt, const __grid_constant__ CUtensorMap tmap_q,
                     const __grid_constant__ CUtensorMap tmap_do, const Attn3Params p) {
  using C = Attn3Cfg<NKT>;
  constexpr int NCH = C::kNCH, NB = C::kNBuf;
#if PEFT_GEMM_DIAG        // cycle counters and the diagnostic A/B flags only in the diagnostics build (make DIAG=1)
  unsigned long long* const dstats = p.stats;
  const int dflags = p.flags;
#else
  constexpr unsigned long long* dstats = nullptr;
  constexpr int dflags = 0;
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sQD = smem;                                          // [2 heads][Q | dO]   (all query rows of the head)
  uint8_t* sT = sQD + 4 * C::kHeadBytes;                        // [2 tiles][K_j | V_j]
  float* sLD = reinterpret_cast<float*>(sT + 4 * C::kTileBytes);   // [2 warpgroups][lse[NKT] | delta[NKT]]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sLD + 4 * NKT);
  uint64_t* qd_full = bars;            // [2]
  uint64_t* qd_empty = bars + 2;       // [2]
  uint64_t* t_full = bars + 4;         // [2]
  uint64_t* t_empty = bars + 6;        // [2]
  uint64_t* sdp_ready = bars + 8;      // [3]
  uint64_t* pds_ready = bars + 11;     // [3]
  uint64_t* out_ready = bars + 14;     // [2]  indexed by tile parity: the warpgroups take the epilogues of alternate tiles when a
                                       //      tile has an odd number of items, so each must see consecutive phases of ITS barrier
  uint64_t* out_free = bars + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int L = p.L, H = p.H;
  const int TK = (L + 127) / 128;
  const int total_heads = p.B * H;
  const int my_heads = (total_heads - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int G = my_heads * TK;
  const int N = G * NCH;
  const int D = H * 64;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_kt); ptx::prefetch_tmap(&tmap_q); ptx::prefetch_tmap(&tmap_do);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&qd_full[i], 1);
      ptx::mbar_init(&qd_empty[i], 1);
      ptx::mbar_init(&t_full[i], 1);
      ptx::mbar_init(&t_empty[i], 1);
    }
    for (int i = 0; i < NB; ++i) {
      ptx::mbar_init(&sdp_ready[i], 1);
      ptx::mbar_init(&pds_ready[i], 128);
    }
    ptx::mbar_init(&out_ready[0], 1);
    ptx::mbar_init(&out_ready[1], 1);
    ptx::mbar_init(out_free, 128);
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc<512>(tmem_slot);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                            // PDL (common.cuh): no global-memory access above this line

  if (warp == 0) {
    // ===================== TMA producer (converged warp, elected lane issues) =====================
    auto prefetch = [&](int g2) {     // L2 prefetch of tile g2's operands (see the dQ kernel)
      if (g2 >= G || !p.prefetch) return;
      const int hl2 = g2 / TK, j2 = g2 % TK;
      const int bh2 = (int)blockIdx.x + hl2 * (int)gridDim.x;
      const int b2 = bh2 / H, h2 = bh2 % H;
      if (ptx::elect_one()) {
        if (j2 == 0) {
#pragma unroll
          for (int r = 0; r < NKT; r += C::kBox) {
            ptx::tma_prefetch_3d(&tmap_q, h2 * 64, r, b2);
            ptx::tma_prefetch_3d(&tmap_do, h2 * 64, r, b2);
          }
        }
        ptx::tma_prefetch_3d(&tmap_kt, D + h2 * 64, j2 * 128, b2);
        ptx::tma_prefetch_3d(&tmap_kt, 2 * D + h2 * 64, j2 * 128, b2);
      }
      __syncwarp();
    };
    prefetch(2);
    prefetch(3);
    for (int g = 0; g < G; ++g) {
      const int hl = g / TK, j = g % TK;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      prefetch(g + 4);
      if (j == 0) {
        const int qdb = hl & 1;
        ptx::mbar_wait(&qd_empty[qdb], ((hl >> 1) & 1) ^ 1);
        uint8_t* sQ = sQD + qdb * 2 * C::kHeadBytes;
        uint8_t* sdO = sQ + C::kHeadBytes;
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&qd_full[qdb], 2 * C::kHeadBytes);
#pragma unroll
          for (int r = 0; r < NKT; r += C::kBox) {
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
    constexpr uint32_t idesc_s64 = ptx::make_idesc_bf16(128, 64);
    constexpr uint32_t idesc_sl = ptx::make_idesc_bf16(128, C::kWLast);
    constexpr uint32_t idesc_o = ptx::make_idesc_bf16_ex(128, 64, 0, 1);
    constexpr uint32_t HI = ptx::kDescHiSw128;
    const uint32_t sT_u32 = ptx::smem_u32(sT), sQD_u32 = ptx::smem_u32(sQD);
    // S^T = K_j Q_c^T and dP^T = V_j dO_c^T of item k into buffer k % 3
    auto issue_ss = [&](int k) {
      const int g = k / NCH, qc = k % NCH;
      const int hl = g / TK, j = g % TK;
      const int qdb = hl & 1, kb = g & 1;
      if (qc == 0) {
        if (j == 0) ptx::mbar_wait(&qd_full[qdb], (hl >> 1) & 1);
        ptx::mbar_wait(&t_full[kb], (g >> 1) & 1);
        ptx::tcgen05_fence_after();
      }
      const uint32_t skj = sT_u32 + kb * 2 * C::kTileBytes;
      const uint32_t svj = skj + C::kTileBytes;
      const uint32_t sq = sQD_u32 + qdb * 2 * C::kHeadBytes + qc * 64 * 128;
      const uint32_t sdo = sq + C::kHeadBytes;
      const uint32_t tS = tmem_base + (uint32_t)((k % NB) * 128);
      const uint32_t idesc = (qc == NCH - 1) ? idesc_sl : idesc_s64;
      if (ptx::elect_one()) {
        const uint32_t k_lo = ptx::desc_lo(skj), q_lo = ptx::desc_lo(sq), v_lo = ptx::desc_lo(svj), do_lo = ptx::desc_lo(sdo);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) ptx::umma_bf16_ss_lohi(tS, k_lo + 2 * kk, q_lo + 2 * kk, HI, idesc, kk != 0);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) ptx::umma_bf16_ss_lohi(tS + 64, v_lo + 2 * kk, do_lo + 2 * kk, HI, idesc, kk != 0);
        ptx::umma_commit(&sdp_ready[k % NB]);
      }
      __syncwarp();
    };
    int next_ss = 0;
    long long w_ss = 0, w_ds = 0, w_free = 0, tc0 = 0;
    const long long tbeg = clock64();
    // K_j / V_j are read by the SS products only: their slot goes back to the producer as soon as the tile's last SS product is
    // in the queue (with the TS products of tile-local item kRel), see the dQ kernel
    constexpr int kRel = (NCH >= 4) ? NCH - 3 : 0;
    const int head_items = TK * NCH;
    auto advance = [&](int done) {     // (a) TMEM buffer free, (b) K_j / V_j slot released, (c) Q / dO slot: head hl - 2 is finished
      while (next_ss < N && next_ss < done + NB && (next_ss / NCH - 2) * NCH + kRel + 1 <= done &&
             (next_ss / head_items - 1) * head_items <= done) {
        if (dstats) tc0 = clock64();
        issue_ss(next_ss);
        if (dstats) w_ss += clock64() - tc0;
        ++next_ss;
      }
    };
    advance(0);
    for (int i = 0; i < N; ++i) {
      const int g = i / NCH, qc = i % NCH;
      const int hl = g / TK, j = g % TK;
      const int qdb = hl & 1, kb = g & 1;
      if (dstats) tc0 = clock64();
      ptx::mbar_wait(&pds_ready[i % NB], (i / NB) & 1);
      if (dstats) { const long long n_ = clock64(); w_ds += n_ - tc0; tc0 = n_; }
      if (qc == 0) ptx::mbar_wait(out_free, (g & 1) ^ 1);     // the previous tile's dK / dV accumulators have been read out
      if (dstats) w_free += clock64() - tc0;
      ptx::tcgen05_fence_after();
      const uint32_t sq = sQD_u32 + qdb * 2 * C::kHeadBytes + qc * 64 * 128;
      const uint32_t sdo = sq + C::kHeadBytes;
      const uint32_t tS = tmem_base + (uint32_t)((i % NB) * 128);
      const uint32_t tDV = tmem_base + (uint32_t)C::kAccCol, tDK = tDV + 64;
      if (ptx::elect_one()) {
        const uint32_t q_lo = ptx::desc_lo(sq), do_lo = ptx::desc_lo(sdo);
        if (C::kWLast == 64 || qc != NCH - 1) {
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const uint32_t acc = (qc | jj) != 0 ? 1u : 0u;
            ptx::umma_bf16_ts_lohi(tDV, tS + 8 * jj, do_lo + jj * (2048 >> 4), HI, idesc_o, acc);
            ptx::umma_bf16_ts_lohi(tDK, tS + 64 + 8 * jj, q_lo + jj * (2048 >> 4), HI, idesc_o, acc);
          }
        } else {
#pragma unroll
          for (int jj = 0; jj < C::kWLast / 16; ++jj) {
            ptx::umma_bf16_ts_lohi(tDV, tS + 8 * jj, do_lo + jj * (2048 >> 4), HI, idesc_o, 1u);
            ptx::umma_bf16_ts_lohi(tDK, tS + 64 + 8 * jj, q_lo + jj * (2048 >> 4), HI, idesc_o, 1u);
          }
        }
        if (qc == kRel) ptx::umma_commit(&t_empty[kb]);
        if (qc == NCH - 1) {
          ptx::umma_commit(&out_ready[g & 1]);
          if (j == TK - 1) ptx::umma_commit(&qd_empty[qdb]);
        }
      }
      __syncwarp();
      advance(i + 1);
    }
    if (dstats && lane == 0) {
      unsigned long long* o = dstats + (256 + blockIdx.x) * 16;
      o[0] = clock64() - tbeg; o[1] = w_ss; o[2] = w_ds; o[3] = w_free;
    }
  } else if (warp >= 4) {
    // ===================== compute warps: warpgroup w takes the items i = w, w + 2, ... =====================
    const int quarter = warp & 3;
    const int wg = (warp - 4) >> 2;
    const int rit = quarter * 32 + lane;
    const float c = p.scale * 1.4426950408889634f;
    const uint32_t tb = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    float* sLse = sLD + wg * 2 * NKT;
    float* sDelta = sLse + NKT;
    long long c_delta = 0, c_wait = 0, c_math = 0, c_epi = 0, tc0 = clock64();
    const long long tbeg = tc0;
    auto lap = [&](long long& acc) { if (dstats) { const long long n_ = clock64(); acc += n_ - tc0; tc0 = n_; } };
    int cur_hl = -1;
    // lse / delta of the NEXT head this warpgroup will meet travel in registers (fetched one head ahead, so their DRAM latency
    // is never exposed): thread rit holds queries rit, rit + 128, rit + 256
    constexpr int kLD = (NKT + 127) / 128;
    const int head_items = TK * NCH;
    float nl[kLD], nd[kLD];
    auto fetch_head = [&](int hl_) {
      const long long base = (long long)((int)blockIdx.x + hl_ * (int)gridDim.x) * L;
#pragma unroll
      for (int k = 0; k < kLD; ++k) {
        const int q = rit + 128 * k;
        nl[k] = q < L ? __ldg(p.lse2 + base + q) : 0.f;
        nd[k] = q < L ? __ldg(p.delta_in + base + q) : 0.f;
      }
    };
    if (wg < N) fetch_head(wg / head_items);
    for (int i = wg; i < N; i += 2) {
      const int g = i / NCH, qc = i % NCH;
      const int hl = g / TK, j = g % TK;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const bool warp_active = (j * 128 + quarter * 32) < L;
      if (hl != cur_hl) {
        // per-head lse / delta of every query (columns of S^T) -> this warpgroup's private copy
        cur_hl = hl;
        ptx::named_bar_sync(2 + wg, 128);            // every thread of the warpgroup is done with the previous head's values
#pragma unroll
        for (int k = 0; k < kLD; ++k) {
          const int q = rit + 128 * k;
          if (q < NKT) { sLse[q] = nl[k]; sDelta[q] = nd[k]; }
        }
        ptx::named_bar_sync(2 + wg, 128);
        // first item of this warpgroup beyond the current head
        const int e = (hl + 1) * head_items;
        const int in = e + ((e ^ wg) & 1);
        if (in < N) fetch_head(in / head_items);
      }
      const int buf = i % NB;
      lap(c_delta);
      ptx::mbar_wait(&sdp_ready[buf], (i / NB) & 1);
      ptx::tcgen05_fence_after();
      lap(c_wait);
      if (warp_active && !(dflags & 2)) {
        const uint32_t tS = tb + (uint32_t)(buf * 128);
        const int qbase = qc * 64;
        const int valid = L - qbase;
        if (C::kWLast == 64 || qc != NCH - 1) {
          if (dflags & 1) {
            a3_dkv_chunk<32>(tS, tS + 64, tS, tS + 64, valid, c, sLse + qbase, sDelta + qbase);
            a3_dkv_chunk<32>(tS + 32, tS + 96, tS + 16, tS + 80, valid - 32, c, sLse + qbase + 32, sDelta + qbase + 32);
          } else {
            a3_dkv_item<64>(tS, valid, c, sLse + qbase, sDelta + qbase);
          }
        } else {
          a3_dkv_chunk<16>(tS, tS + 64, tS, tS + 64, valid, c, sLse + qbase, sDelta + qbase);
        }
        ptx::tmem_st_wait();
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(&pds_ready[buf]);
      lap(c_math);
      if (qc == NCH - 1) {
        // ---- epilogue of the key tile (by the warpgroup that took its last item): dV_j and dK_j * scale -> bf16 -> global
        const int b = bh / H, h = bh % H;
        const int row = j * 128 + rit;
        ptx::mbar_wait(&out_ready[g & 1], (g >> 1) & 1);
        ptx::tcgen05_fence_after();
        // four 32-column quarters (dV | dK): load, pack, store 64 bytes — few live registers; out_free after the last load
        __nv_bfloat16* dst = p.dqkv + ((long long)b * L + row) * p.lddqkv + h * 64;
#pragma unroll
        for (int qt = 0; qt < 4; ++qt) {
          uint32_t r[32], pk[16];
          if (warp_active) {
            tm_ld32(tb + C::kAccCol + 32 * qt, r);
            ptx::tmem_ld_wait();
          }
          if (qt == 3) {
            ptx::tcgen05_fence_before();
            ptx::mbar_arrive(out_free);
          }
          if (warp_active && row < L) {
            const float f = qt < 2 ? 1.0f : p.scale;
#pragma unroll
            for (int q = 0; q < 16; ++q) pk[q] = ptx::pack_bf16x2(__uint_as_float(r[2 * q]) * f, __uint_as_float(r[2 * q + 1]) * f);
            uint4* d4 = reinterpret_cast<uint4*>(dst + (qt < 2 ? 2 * D : D) + 32 * (qt & 1));
#pragma unroll
            for (int q = 0; q < 4; ++q) d4[q] = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
          }
        }
        lap(c_epi);
      }
    }
    if (dstats && (warp & 3) == 0 && lane == 0) {
      unsigned long long* o = dstats + (256 + blockIdx.x) * 16 + 4 + 6 * wg;
      o[0] = clock64() - tbeg; o[1] = c_delta; o[2] = c_wait; o[3] = c_math; o[4] = c_epi;
    }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  pdl_trigger();                                         // PDL: late on purpose (see gemm2_sm100.cuh)
  if (warp == 2) ptx::tmem_dealloc<512>(tmem_base);
}

}  // namespace peft

// ---------------------------------------------------------------------------------------------------------------------
// Micro-probe (diagnostics, peft_debug_mma_probe): cycles per tcgen05.mma (M = 128, K = 16, bf16) for chains of `reps`
// instructions that accumulate into `nacc` different TMEM accumulators in rotation (nacc = 1: every instruction depends on
// its predecessor).  ts = 1: A operand from TMEM (the P.V / dS.K form), else both operands from shared memory.
// ---------------------------------------------------------------------------------------------------------------------
namespace peft {
template <int NACC>
__global__ void __launch_bounds__(128, 1) mma_probe_kernel(int n, int ts, int reps, unsigned long long* out) {
  extern __shared__ uint8_t probe_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(probe_smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 2 * 32768 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
  if (threadIdx.x < 32) ptx::tmem_alloc<512>(&slot);
  ptx::fence_proxy_async();
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x < 32) {
    const uint32_t idesc = ts ? ptx::make_idesc_bf16_ex(128, n, 0, 1) : ptx::make_idesc_bf16(128, n);
    const uint32_t a_lo = ptx::desc_lo(ptx::smem_u32(smem)), b_lo = ptx::desc_lo(ptx::smem_u32(smem + 32768));
    const long long t0 = clock64();
    for (int r = 0; r < reps; r += 8) {          // 8 instructions per trip, everything but `n` a compile-time constant
      if (ptx::elect_one()) {
        if (ts) {
#pragma unroll
          for (int u = 0; u < 8; ++u)
            ptx::umma_bf16_ts_lohi(tm + 256 + (uint32_t)((u % NACC) * n), tm + 8 * u, b_lo, ptx::kDescHiSw128, idesc, 1u);
        } else {
#pragma unroll
          for (int u = 0; u < 8; ++u)
            ptx::umma_bf16_ss_lohi(tm + 256 + (uint32_t)((u % NACC) * n), a_lo + 2 * (u & 3), b_lo + 2 * (u & 3), ptx::kDescHiSw128, idesc, 1u);
        }
      }
      __syncwarp();
    }
    const long long t1 = clock64();
    if (ptx::elect_one()) ptx::umma_commit(&bar);
    __syncwarp();
    ptx::mbar_wait(&bar, 0);
    const long long t2 = clock64();
    if (threadIdx.x == 0) { out[0] = (unsigned long long)(t1 - t0); out[1] = (unsigned long long)(t2 - t0); }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc<512>(tm);
}
}  // namespace peft
