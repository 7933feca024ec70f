// tcgen05 / TMEM / TMA multi-head attention for the ViT sequence lengths (L <= 272, head_dim 64): forward.
//
// replaces: the (B*H, L, L) score / softmax / PV chain of FS/kadaptation.py:202-205,223-235 and FS/model.py:907-918,
// 990-1002.  One persistent CTA per SM walks over (batch, head) pairs; the whole K and V of a
// head ([NK x 64] bf16 each, NK = L padded to 16) sit in shared memory, fetched by 3-D TMA straight out of the packed
// projection buffer qkv[B, L, 3D] (rows >= L are zero-filled by the TMA unit, so no padding copy exists anywhere).
// Per 128-query tile:
//     S = Q K^T        tcgen05.mma SS (A = Q tile, B = K, both K-major), fp32 in TMEM            [128 x NK]
//     P = exp2(S c - m c), row sums                 8 warps: TMEM lane = query row, two warpgroups split the columns;
//                                                   P (bf16) is written back IN PLACE into the S columns
//     O = P V          tcgen05.mma TS (A = P from TMEM, B = V as MN-major smem operand)           [128 x 64]
//     o = O / rowsum -> bf16 -> swizzled smem -> 3-D TMA store (rows >= L clipped), lse2 = m c + log2(rowsum)
// With NK <= 256 two S/P/O buffers live in TMEM so the tensor core works on tile g+1 while the warps exponentiate tile g.
#pragma once
#include "gemm2_sm100.cuh"

namespace peft {

constexpr int kA2Threads = 384;   // warp 0 TMA, warp 1 MMA, warp 2 TMEM alloc, warps 4..11 compute

namespace ptx {
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// L2 prefetch of a 3-D box (no shared memory, no barrier).  Built to test whether the attention kernels are bound by the latency
// of their operand loads: they are not (no gain with every box prefetched four tiles ahead); only the pipelined
// backward (attention3_bwd_sm100.cuh) still carries the switch (PEFT_ATTN_PREFETCH=1).
__device__ __forceinline__ void tma_prefetch_3d(const CUtensorMap* tm, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(c0),
               "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* tm, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// ---- descriptor-by-halves forms: {lo = (addr >> 4) | LBO field 1, hi = SBO 1024 B | version 1 | SWIZZLE_128B}; K-major tiles
// advance lo by 2 per 16-element K step, MN-major tiles by 128 (2 KB) — 32-bit uniform-datapath arithmetic only
constexpr uint32_t kDescHiSw128 = (1024u >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr) { return ((smem_addr & 0x3FFFFu) >> 4) | (1u << 16); }
__device__ __forceinline__ void umma_bf16_ss_lohi(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .b64 da, db;\n"
      "setp.ne.b32 p, %5, 0;\n"
      "mov.b64 da, {%1, %3};\n"
      "mov.b64 db, {%2, %3};\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_ts_lohi(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .b64 db;\n"
      "setp.ne.b32 p, %5, 0;\n"
      "mov.b64 db, {%2, %3};\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %4, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]),
               "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// MN-major operand tile in shared memory: rows = K index (128-byte rows of 64 contiguous MN elements), SWIZZLE_128B,
// 8-row groups 1024 B apart (SBO); LBO (distance between 64-element MN atoms) is unused for MN extent 64.
__device__ __forceinline__ uint64_t make_smem_desc_sw128_mn(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
// kind::f16 instruction descriptor with selectable operand majors (bit 15: A is MN-major, bit 16: B is MN-major)
__host__ __device__ constexpr uint32_t make_idesc_bf16_ex(uint32_t M, uint32_t N, uint32_t a_mn, uint32_t b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
}  // namespace ptx

struct Attn2Params {
  float* lse2;      // [B*H, L]
  const float* delta_in;   // backward (dK/dV kernel): [B*H, L]
  float* delta_out;        // backward (dQ kernel): [B*H, L]
  int B, H, L;
  float scale;
  unsigned long long* stats;   // diagnostics (PEFT_ATTN_STATS=1): cycles per phase of the first softmax warp of each CTA
};


// ---- register-array TMEM helpers (pointer form: with full unrolling the indices are compile-time constants)
#define PEFT_R8(r, o) "=r"(r[o]), "=r"(r[o + 1]), "=r"(r[o + 2]), "=r"(r[o + 3]), "=r"(r[o + 4]), "=r"(r[o + 5]), "=r"(r[o + 6]), "=r"(r[o + 7])
#define PEFT_W8(r, o) "r"(r[o]), "r"(r[o + 1]), "r"(r[o + 2]), "r"(r[o + 3]), "r"(r[o + 4]), "r"(r[o + 5]), "r"(r[o + 6]), "r"(r[o + 7])
__device__ __forceinline__ void tm_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : PEFT_R8(r, 0), PEFT_R8(r, 8), PEFT_R8(r, 16), PEFT_R8(r, 24)
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tm_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : PEFT_R8(r, 0), PEFT_R8(r, 8)
      : "r"(taddr)
      : "memory");
}
template <int W>
__device__ __forceinline__ void tm_ld(uint32_t taddr, uint32_t* r) {
  if constexpr (W == 32) tm_ld32(taddr, r);
  else tm_ld16(taddr, r);
}
__device__ __forceinline__ void tm_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      PEFT_W8(r, 0), PEFT_W8(r, 8)
      : "memory");
}
__device__ __forceinline__ void tm_st8(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), PEFT_W8(r, 0)
               : "memory");
}
template <int W>   // W = number of fp32 source columns (32 or 16) -> W/2 packed columns
__device__ __forceinline__ void tm_st_packed(uint32_t taddr, const uint32_t* r) {
  if constexpr (W == 32) tm_st16(taddr, r);
  else tm_st8(taddr, r);
}

// =====================================================================================================================
// forward
// =====================================================================================================================
template <int NK>
struct Attn2Cfg {
  static_assert(NK % 16 == 0 && NK >= 64 && NK <= 272, "padded key count");
  static constexpr int kNCh = (NK + 31) / 32;                          // 32-column chunks (the last may be 16 wide)
  static constexpr int kOAlias = ((NK / 2 + 31) / 32) * 32;           // P occupies [0, NK/2); O aliased above it if it fits
  static constexpr int kOOff = (kOAlias + 64 <= NK) ? kOAlias : NK;
  static constexpr int kBufCols = (kOOff + 64 > NK) ? (kOOff + 64) : NK;
  static constexpr int kBufs = (2 * kBufCols <= 512) ? 2 : 1;
  static constexpr int kTmemCols = (kBufs * kBufCols <= 128) ? 128 : ((kBufs * kBufCols <= 256) ? 256 : 512);
  static constexpr int kKVBox = (NK <= 256) ? NK : NK / 2;            // rows per TMA box of K / V
  static constexpr int kKVBytes = NK * 128;
  static constexpr int kQBytes = 128 * 128;
  // smem: K,V x2 heads | Q x2 tiles | O staging x2 (one per warpgroup) | barriers
  static constexpr int kSmemBytes = 4 * kKVBytes + 2 * kQBytes + 2 * kQBytes + 256 + 1024;
};

// Forward kernel.  Warpgroup w (warps 4+4w .. 7+4w) owns the tiles g = w, w+2, ...: one thread per query row walks all
// NK columns (pass 1: row max, pass 2: exp2 / row sum / bf16 P written in place), so no cross-thread reduction exists
// and the two warpgroups run out of phase, which keeps the MUFU pipe and the tensor core busy at the same time.
template <int NK>
__global__ void __launch_bounds__(kA2Threads, 1)
attn2_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                 const __grid_constant__ CUtensorMap tmap_o, const Attn2Params p) {
  using C = Attn2Cfg<NK>;
#if PEFT_GEMM_DIAG        // phase counters only in the diagnostics build (make DIAG=1), see gemm2_sm100.cuh
  unsigned long long* const dstats = p.stats;
#else
  constexpr unsigned long long* dstats = nullptr;
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sKV = smem;                                   // [2 heads][K | V]
  uint8_t* sQ = sKV + 4 * C::kKVBytes;                   // [2 tiles]
  uint8_t* sO = sQ + 2 * C::kQBytes;                     // [2 warpgroups] staging for the TMA store
  uint64_t* bars = reinterpret_cast<uint64_t*>(sO + 2 * C::kQBytes);
  uint64_t* kv_full = bars;            // [2]
  uint64_t* kv_empty = bars + 2;       // [2]
  uint64_t* q_full = bars + 4;         // [2]
  uint64_t* q_empty = bars + 6;        // [2]
  uint64_t* s_ready = bars + 8;        // [2]
  uint64_t* p_ready = bars + 10;       // [2]
  uint64_t* o_ready = bars + 12;       // [2]
  uint64_t* buf_free = bars + 14;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int L = p.L, H = p.H;
  const int TQ = (L + 127) / 128;
  const int total_heads = p.B * H;
  const int my_heads = (total_heads - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int G = my_heads * TQ;
  const int D = H * 64;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_q);
    ptx::prefetch_tmap(&tmap_kv);
    ptx::prefetch_tmap(&tmap_o);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&kv_full[i], 1);
      ptx::mbar_init(&kv_empty[i], 1);
      ptx::mbar_init(&q_full[i], 1);
      ptx::mbar_init(&q_empty[i], 1);
      ptx::mbar_init(&s_ready[i], 1);
      ptx::mbar_init(&p_ready[i], 128);
      ptx::mbar_init(&o_ready[i], 1);
      ptx::mbar_init(&buf_free[i], 128);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc<C::kTmemCols>(tmem_slot);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                            // no global-memory access above this line

  if (warp == 0) {
    // ===================== TMA producer (converged warp, elected lane issues) =====================
    for (int g = 0; g < G; ++g) {
      const int hl = g / TQ, t = g % TQ;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      if (t == 0) {
        const int kvb = hl & 1;
        ptx::mbar_wait(&kv_empty[kvb], ((hl >> 1) & 1) ^ 1);
        uint8_t* sK = sKV + kvb * 2 * C::kKVBytes;
        uint8_t* sV = sK + C::kKVBytes;
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&kv_full[kvb], 2 * C::kKVBytes);
#pragma unroll
          for (int r = 0; r < NK; r += C::kKVBox) {
            ptx::tma_load_3d(sK + r * 128, &tmap_kv, &kv_full[kvb], D + h * 64, r, b);
            ptx::tma_load_3d(sV + r * 128, &tmap_kv, &kv_full[kvb], 2 * D + h * 64, r, b);
          }
        }
        __syncwarp();
      }
      const int qb = g & 1;
      ptx::mbar_wait(&q_empty[qb], ((g >> 1) & 1) ^ 1);
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(&q_full[qb], C::kQBytes);
        ptx::tma_load_3d(sQ + qb * C::kQBytes, &tmap_q, &q_full[qb], h * 64, t * 128, b);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The warp runs converged; only the tcgen05.mma / commit instructions sit under elect.sync, so that their descriptors are
    // warp-uniform values in uniform registers (a loop inside `if (lane == 0)` costs an R2UR waterfall of ~25 dependent
    // scalar instructions per MMA, several times the 32-clk cost of a 128 x 64 x 16 P.V step).
    {
      constexpr uint32_t idesc_pv = ptx::make_idesc_bf16_ex(128, 64, 0, 1);
      const uint32_t sKV_u32 = ptx::smem_u32(sKV), sQ_u32 = ptx::smem_u32(sQ);
      auto issue_pv = [&](int g) {
        const int hl = g / TQ, t = g % TQ;
        const int kvb = hl & 1;
        const int buf = g % C::kBufs;
        ptx::mbar_wait(&p_ready[g & 1], (g >> 1) & 1);
        ptx::tcgen05_fence_after();
        const uint32_t sV = sKV_u32 + kvb * 2 * C::kKVBytes + C::kKVBytes;
        const uint32_t tbuf = tmem_base + (uint32_t)(buf * C::kBufCols);
        if (ptx::elect_one()) {
          const uint32_t v_lo = ptx::desc_lo(sV);
#pragma unroll
          for (int j = 0; j < NK / 16; ++j)
            ptx::umma_bf16_ts_lohi(tbuf + C::kOOff, tbuf + 8 * j, v_lo + j * (2048 >> 4), ptx::kDescHiSw128, idesc_pv, j != 0 ? 1u : 0u);
          ptx::umma_commit(&o_ready[g & 1]);
          if (t == TQ - 1) ptx::umma_commit(&kv_empty[kvb]);
        }
        __syncwarp();
      };
      for (int g = 0; g < G; ++g) {
        const int hl = g / TQ, t = g % TQ;
        const int kvb = hl & 1, qb = g & 1, buf = g % C::kBufs;
        if (t == 0) ptx::mbar_wait(&kv_full[kvb], (hl >> 1) & 1);
        ptx::mbar_wait(&q_full[qb], (g >> 1) & 1);
        // barriers facing the warpgroups are indexed by the warpgroup slot g & 1 (use count g >> 1); the TMEM buffer is
        // g % kBufs, so with a single buffer the previous user of the buffer is tile g - 1
        if constexpr (C::kBufs == 2) {
          ptx::mbar_wait(&buf_free[g & 1], ((g >> 1) & 1) ^ 1);
        } else {
          if (g >= 1) ptx::mbar_wait(&buf_free[(g - 1) & 1], ((g - 1) >> 1) & 1);
        }
        ptx::tcgen05_fence_after();
        const uint32_t sq = sQ_u32 + qb * C::kQBytes;
        const uint32_t sk = sKV_u32 + kvb * 2 * C::kKVBytes;
        const uint32_t tbuf = tmem_base + (uint32_t)(buf * C::kBufCols);
        if (ptx::elect_one()) {
          const uint32_t q_lo = ptx::desc_lo(sq), k_lo = ptx::desc_lo(sk);
          if constexpr (NK <= 256) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(128, NK);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_bf16_ss_lohi(tbuf, q_lo + 2 * k, k_lo + 2 * k, ptx::kDescHiSw128, idesc, k != 0);
          } else {
            constexpr int N0 = 144, N1 = NK - 144;
            constexpr uint32_t idesc0 = ptx::make_idesc_bf16(128, N0), idesc1 = ptx::make_idesc_bf16(128, N1);
            const uint32_t k1_lo = ptx::desc_lo(sk + N0 * 128);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_bf16_ss_lohi(tbuf, q_lo + 2 * k, k_lo + 2 * k, ptx::kDescHiSw128, idesc0, k != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_bf16_ss_lohi(tbuf + N0, q_lo + 2 * k, k1_lo + 2 * k, ptx::kDescHiSw128, idesc1, k != 0);
          }
          ptx::umma_commit(&q_empty[qb]);
          ptx::umma_commit(&s_ready[g & 1]);
        }
        __syncwarp();
        if constexpr (C::kBufs == 2) {
          if (g >= 1) issue_pv(g - 1);
        } else {
          issue_pv(g);
        }
      }
      if constexpr (C::kBufs == 2) {
        if (G >= 1) issue_pv(G - 1);
      }
    }
  } else if (warp >= 4) {
    // ===================== softmax / epilogue warpgroups =====================
    const int cw = warp - 4;
    const int quarter = warp & 3;          // TMEM lane quarter
    const int wg = cw >> 2;
    const int rit = quarter * 32 + lane;   // row in tile == TMEM lane
    const float c = p.scale * 1.4426950408889634f;
    const bool issuer = ((cw & 3) == 0) && lane == 0;
    uint8_t* stage = sO + wg * C::kQBytes;
    constexpr int NF = NK / 32;
    constexpr bool kTail = (NK % 32) != 0;
    long long tS = 0, tP1 = 0, tP2 = 0, tO = 0, tE = 0, tc0 = clock64(), tbeg = tc0;
    auto lap = [&](long long& acc) { const long long n = clock64(); acc += n - tc0; tc0 = n; };
    for (int g = wg; g < G; g += 2) {
      const int hl = g / TQ, t = g % TQ;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      const int buf = g % C::kBufs;
      const int row = t * 128 + rit;
      const bool warp_active = (t * 128 + quarter * 32) < L;
      const uint32_t tb = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + (uint32_t)(buf * C::kBufCols);
      ptx::mbar_wait(&s_ready[wg], (g >> 1) & 1);
      ptx::tcgen05_fence_after();
      if (dstats) lap(tS);
      float mx = -INFINITY, sum = 0.f, mc = 0.f;
      if (warp_active) {
        uint32_t r[2][32];
        // ---- pass 1: row maximum
        tm_ld32(tb, r[0]);
#pragma unroll
        for (int ci = 0; ci < NF; ++ci) {
          ptx::tmem_ld_wait();
          if (ci + 1 < NF) tm_ld32(tb + (ci + 1) * 32, r[(ci + 1) & 1]);
          else if (kTail) tm_ld16(tb + NF * 32, r[(ci + 1) & 1]);
          const int c0 = ci * 32;
          if (c0 + 32 <= L) {
#pragma unroll
            for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(r[ci & 1][j]));
          } else if (c0 < L) {
#pragma unroll
            for (int j = 0; j < 32; ++j) mx = fmaxf(mx, (c0 + j < L) ? __uint_as_float(r[ci & 1][j]) : -INFINITY);
          }
        }
        if constexpr (kTail) {
          ptx::tmem_ld_wait();
          constexpr int c0 = NF * 32;
          if (c0 < L) {
#pragma unroll
            for (int j = 0; j < 16; ++j) mx = fmaxf(mx, (c0 + j < L) ? __uint_as_float(r[NF & 1][j]) : -INFINITY);
          }
        }
        mc = mx * c;
        if (dstats) lap(tP1);
        // ---- pass 2: P = exp2(S c - m c) -> bf16 in place (ascending: writes only touch columns already consumed)
        tm_ld32(tb, r[0]);
#pragma unroll
        for (int ci = 0; ci < NF; ++ci) {
          ptx::tmem_ld_wait();
          if (ci + 1 < NF) tm_ld32(tb + (ci + 1) * 32, r[(ci + 1) & 1]);
          else if (kTail) tm_ld16(tb + NF * 32, r[(ci + 1) & 1]);
          const int c0 = ci * 32;
          uint32_t pk[16];
          if (c0 + 32 <= L) {
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float p0 = ptx::ex2_approx(fmaf(__uint_as_float(r[ci & 1][j]), c, -mc));
              const float p1 = ptx::ex2_approx(fmaf(__uint_as_float(r[ci & 1][j + 1]), c, -mc));
              sum += p0 + p1;
              pk[j >> 1] = ptx::pack_bf16x2(p0, p1);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float p0 = (c0 + j < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[ci & 1][j]), c, -mc)) : 0.f;
              const float p1 = (c0 + j + 1 < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[ci & 1][j + 1]), c, -mc)) : 0.f;
              sum += p0 + p1;
              pk[j >> 1] = ptx::pack_bf16x2(p0, p1);
            }
          }
          tm_st16(tb + (c0 >> 1), pk);
        }
        if constexpr (kTail) {
          ptx::tmem_ld_wait();
          constexpr int c0 = NF * 32;
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 16; j += 2) {
            const float p0 = (c0 + j < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[NF & 1][j]), c, -mc)) : 0.f;
            const float p1 = (c0 + j + 1 < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[NF & 1][j + 1]), c, -mc)) : 0.f;
            sum += p0 + p1;
            pk[j >> 1] = ptx::pack_bf16x2(p0, p1);
          }
          tm_st8(tb + (c0 >> 1), pk);
        }
        ptx::tmem_st_wait();
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(&p_ready[wg]);
      if (dstats) lap(tP2);
      // ---- epilogue: O / rowsum -> bf16 -> staging -> TMA store
      ptx::mbar_wait(&o_ready[wg], (g >> 1) & 1);
      ptx::tcgen05_fence_after();
      if (dstats) lap(tO);
      uint32_t o[64];
      if (warp_active) {
        tm_ld32(tb + C::kOOff, o);
        tm_ld32(tb + C::kOOff + 32, o + 32);
        ptx::tmem_ld_wait();
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(&buf_free[wg]);
      if (issuer) ptx::bulk_wait_read<0>();            // this warpgroup's previous store has drained the staging tile
      ptx::named_bar_sync(1 + wg, 128);
      if (warp_active) {
        const float inv = 1.0f / sum;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint4 u;
          u.x = ptx::pack_bf16x2(__uint_as_float(o[8 * j]) * inv, __uint_as_float(o[8 * j + 1]) * inv);
          u.y = ptx::pack_bf16x2(__uint_as_float(o[8 * j + 2]) * inv, __uint_as_float(o[8 * j + 3]) * inv);
          u.z = ptx::pack_bf16x2(__uint_as_float(o[8 * j + 4]) * inv, __uint_as_float(o[8 * j + 5]) * inv);
          u.w = ptx::pack_bf16x2(__uint_as_float(o[8 * j + 6]) * inv, __uint_as_float(o[8 * j + 7]) * inv);
          *reinterpret_cast<uint4*>(stage + rit * 128 + ((j ^ (rit & 7)) << 4)) = u;
        }
        if (row < L) p.lse2[(long long)bh * L + row] = mc + log2f(sum);
      }
      ptx::fence_proxy_async();
      ptx::named_bar_sync(1 + wg, 128);
      if (issuer) {
        ptx::tma_store_3d(&tmap_o, stage, h * 64, t * 128, b);
        ptx::bulk_commit();
      }
      if (dstats) lap(tE);
    }
    if (issuer) ptx::bulk_wait<0>();
    if (dstats && (cw & 3) == 0 && lane == 0) {
      unsigned long long* o = dstats + (blockIdx.x * 2 + wg) * 8;
      o[0] = tS; o[1] = tP1; o[2] = tP2; o[3] = tO; o[4] = tE; o[5] = clock64() - tbeg;
    }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  pdl_trigger();                                         // PDL: late on purpose (see gemm2_sm100.cuh)
  if (warp == 2) ptx::tmem_dealloc<C::kTmemCols>(tmem_base);
}

}  // namespace peft
