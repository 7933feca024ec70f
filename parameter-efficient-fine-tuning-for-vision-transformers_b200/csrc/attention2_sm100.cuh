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
};

template <int NK>
struct Attn2Cfg {
  static_assert(NK % 16 == 0 && NK >= 64 && NK <= 272, "padded key count");
  static constexpr int kCA = ((NK / 2 + 15) / 16) * 16;            // columns of warpgroup A: [0, kCA); B: [kCA, NK)
  static constexpr int kPB = NK - (NK - kCA) / 2;                  // first TMEM column of warpgroup B's P block
  static constexpr int kOAlias = ((kCA / 2 + 31) / 32) * 32;       // O aliased between the two P blocks when it fits
  static constexpr bool kAliasOK = (kOAlias + 64 <= kPB);
  static constexpr int kOOff = kAliasOK ? kOAlias : NK;
  static constexpr int kBufCols = (kOOff + 64 > NK) ? (kOOff + 64) : NK;
  static constexpr int kBufs = (2 * kBufCols <= 512) ? 2 : 1;
  static constexpr int kTmemCols = (kBufs * kBufCols <= 128) ? 128 : ((kBufs * kBufCols <= 256) ? 256 : 512);
  static constexpr int kKVBox = (NK <= 256) ? NK : NK / 2;         // rows per TMA box of K / V
  static constexpr int kKVBytes = NK * 128;
  static constexpr int kQBytes = 128 * 128;
  static_assert((NK - kCA) % 32 == 0, "warpgroup B processes 32-column chunks");
  // smem: K,V x2 heads | Q x2 tiles | O staging x2 | reductions | barriers
  static constexpr int kSmemBytes = 4 * kKVBytes + 2 * kQBytes + 2 * kQBytes + 4 * 128 * 4 + 256 + 1024;
};

// TMEM column of the P block (A operand of the PV product) holding keys [16 j, 16 j + 16)
template <int NK>
__device__ __forceinline__ uint32_t p_col_of_kstep(int j) {
  using C = Attn2Cfg<NK>;
  const int key = 16 * j;
  return key < C::kCA ? (uint32_t)(key >> 1) : (uint32_t)(NK - ((NK - key) >> 1));
}

template <int NK>
__global__ void __launch_bounds__(kA2Threads, 1)
attn2_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                 const __grid_constant__ CUtensorMap tmap_o, const Attn2Params p) {
  using C = Attn2Cfg<NK>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sKV = smem;                                   // [2 heads][K | V]
  uint8_t* sQ = sKV + 4 * C::kKVBytes;                   // [2 tiles]
  uint8_t* sO = sQ + 2 * C::kQBytes;                     // [2] staging for the TMA store
  float* red_max = reinterpret_cast<float*>(sO + 2 * C::kQBytes);   // [2][128]
  float* red_sum = red_max + 256;                                   // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(red_sum + 256);
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
      ptx::mbar_init(&p_ready[i], 256);
      ptx::mbar_init(&o_ready[i], 1);
      ptx::mbar_init(&buf_free[i], 256);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc<C::kTmemCols>(tmem_slot);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      for (int g = 0; g < G; ++g) {
        const int hl = g / TQ, t = g % TQ;
        const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
        const int b = bh / H, h = bh % H;
        if (t == 0) {
          const int kvb = hl & 1;
          ptx::mbar_wait(&kv_empty[kvb], ((hl >> 1) & 1) ^ 1);
          uint8_t* sK = sKV + kvb * 2 * C::kKVBytes;
          uint8_t* sV = sK + C::kKVBytes;
          ptx::mbar_expect_tx(&kv_full[kvb], 2 * C::kKVBytes);
#pragma unroll
          for (int r = 0; r < NK; r += C::kKVBox) {
            ptx::tma_load_3d(sK + r * 128, &tmap_kv, &kv_full[kvb], D + h * 64, r, b);
            ptx::tma_load_3d(sV + r * 128, &tmap_kv, &kv_full[kvb], 2 * D + h * 64, r, b);
          }
        }
        const int qb = g & 1;
        ptx::mbar_wait(&q_empty[qb], ((g >> 1) & 1) ^ 1);
        ptx::mbar_expect_tx(&q_full[qb], C::kQBytes);
        ptx::tma_load_3d(sQ + qb * C::kQBytes, &tmap_q, &q_full[qb], h * 64, t * 128, b);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc_pv = ptx::make_idesc_bf16_ex(128, 64, 0, 1);
      auto issue_pv = [&](int g) {
        const int hl = g / TQ, t = g % TQ;
        const int kvb = hl & 1;
        const int buf = g % C::kBufs;
        ptx::mbar_wait(&p_ready[buf], (g / C::kBufs) & 1);
        ptx::tcgen05_fence_after();
        const uint32_t sV = ptx::smem_u32(sKV + kvb * 2 * C::kKVBytes + C::kKVBytes);
        const uint32_t tbuf = tmem_base + (uint32_t)(buf * C::kBufCols);
#pragma unroll
        for (int j = 0; j < NK / 16; ++j)
          ptx::umma_bf16_ts(tbuf + C::kOOff, tbuf + p_col_of_kstep<NK>(j), ptx::make_smem_desc_sw128_mn(sV + j * 2048),
                            idesc_pv, j != 0 ? 1u : 0u);
        ptx::umma_commit(&o_ready[buf]);
        if (t == TQ - 1) ptx::umma_commit(&kv_empty[kvb]);
      };
      for (int g = 0; g < G; ++g) {
        const int hl = g / TQ, t = g % TQ;
        const int kvb = hl & 1, qb = g & 1, buf = g % C::kBufs;
        if (t == 0) ptx::mbar_wait(&kv_full[kvb], (hl >> 1) & 1);
        ptx::mbar_wait(&q_full[qb], (g >> 1) & 1);
        ptx::mbar_wait(&buf_free[buf], ((g / C::kBufs) & 1) ^ 1);
        ptx::tcgen05_fence_after();
        const uint32_t sq = ptx::smem_u32(sQ + qb * C::kQBytes);
        const uint32_t sk = ptx::smem_u32(sKV + kvb * 2 * C::kKVBytes);
        const uint32_t tbuf = tmem_base + (uint32_t)(buf * C::kBufCols);
        if constexpr (NK <= 256) {
          constexpr uint32_t idesc = ptx::make_idesc_bf16(128, NK);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            ptx::umma_bf16(tbuf, ptx::make_smem_desc_sw128(sq) + 2 * k, ptx::make_smem_desc_sw128(sk) + 2 * k, idesc, k != 0);
        } else {
          constexpr int N0 = 144, N1 = NK - 144;
          constexpr uint32_t idesc0 = ptx::make_idesc_bf16(128, N0), idesc1 = ptx::make_idesc_bf16(128, N1);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            ptx::umma_bf16(tbuf, ptx::make_smem_desc_sw128(sq) + 2 * k, ptx::make_smem_desc_sw128(sk) + 2 * k, idesc0, k != 0);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            ptx::umma_bf16(tbuf + N0, ptx::make_smem_desc_sw128(sq) + 2 * k, ptx::make_smem_desc_sw128(sk + N0 * 128) + 2 * k,
                           idesc1, k != 0);
        }
        ptx::umma_commit(&q_empty[qb]);
        ptx::umma_commit(&s_ready[buf]);
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
    // ===================== softmax / epilogue warps =====================
    const int cw = warp - 4;
    const int quarter = warp & 3;          // TMEM lane quarter
    const int wg = cw >> 2;                // 0: columns [0, kCA)   1: columns [kCA, NK)
    const int rit = quarter * 32 + lane;   // row in tile == TMEM lane
    const float c = p.scale * 1.4426950408889634f;
    const bool issuer = (threadIdx.x == 128);
    for (int g = 0; g < G; ++g) {
      const int hl = g / TQ, t = g % TQ;
      const int bh = (int)blockIdx.x + hl * (int)gridDim.x;
      const int b = bh / H, h = bh % H;
      const int buf = g % C::kBufs;
      const int row = t * 128 + rit;
      const bool warp_active = (t * 128 + quarter * 32) < L;
      const uint32_t tbuf = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + (uint32_t)(buf * C::kBufCols);
      ptx::mbar_wait(&s_ready[buf], (g / C::kBufs) & 1);
      ptx::tcgen05_fence_after();
      // ---- pass 1: row maximum over this warpgroup's valid columns
      float lmax = -INFINITY;
      if (warp_active) {
        if (wg == 0) {
#pragma unroll 1
          for (int c0 = 0; c0 + 32 <= C::kCA; c0 += 32) {
            uint32_t r[32];
            ptx::tmem_ld_32x32b_x32(tbuf + c0, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) lmax = fmaxf(lmax, (c0 + j < L) ? __uint_as_float(r[j]) : -INFINITY);
          }
          if constexpr (C::kCA % 32 != 0) {
            constexpr int c0 = (C::kCA / 32) * 32;
            uint32_t r[16];
            ptx::tmem_ld_32x32b_x16(tbuf + c0, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) lmax = fmaxf(lmax, (c0 + j < L) ? __uint_as_float(r[j]) : -INFINITY);
          }
        } else {
#pragma unroll 1
          for (int c0 = C::kCA; c0 < NK; c0 += 32) {
            uint32_t r[32];
            ptx::tmem_ld_32x32b_x32(tbuf + c0, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) lmax = fmaxf(lmax, (c0 + j < L) ? __uint_as_float(r[j]) : -INFINITY);
          }
        }
      }
      red_max[wg * 128 + rit] = lmax;
      if (issuer) ptx::bulk_wait_read<1>();            // the staging buffer of tile g-2 has been drained
      ptx::named_bar_sync(1, 256);
      const float m = fmaxf(red_max[rit], red_max[128 + rit]);
      const float mc = m * c;
      // ---- pass 2: P = exp2(S c - m c) -> bf16, written in place; row sum
      float lsum = 0.f;
      if (warp_active) {
        if (wg == 0) {
#pragma unroll 1
          for (int c0 = 0; c0 + 32 <= C::kCA; c0 += 32) {
            uint32_t r[32], pk[16];
            ptx::tmem_ld_32x32b_x32(tbuf + c0, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float p0 = (c0 + j < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[j]), c, -mc)) : 0.f;
              const float p1 = (c0 + j + 1 < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[j + 1]), c, -mc)) : 0.f;
              lsum += p0 + p1;
              pk[j >> 1] = ptx::pack_bf16x2(p0, p1);
            }
            ptx::tmem_st_32x32b_x16(tbuf + (c0 >> 1), pk);
          }
          if constexpr (C::kCA % 32 != 0) {
            constexpr int c0 = (C::kCA / 32) * 32;
            uint32_t r[16], pk[8];
            ptx::tmem_ld_32x32b_x16(tbuf + c0, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
              const float p0 = (c0 + j < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[j]), c, -mc)) : 0.f;
              const float p1 = (c0 + j + 1 < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[j + 1]), c, -mc)) : 0.f;
              lsum += p0 + p1;
              pk[j >> 1] = ptx::pack_bf16x2(p0, p1);
            }
            ptx::tmem_st_32x32b_x8(tbuf + (c0 >> 1), pk);
          }
        } else {
          // descending, so that the P block (top of this warpgroup's range) only overwrites columns already consumed
#pragma unroll 1
          for (int c0 = NK - 32; c0 >= C::kCA; c0 -= 32) {
            uint32_t r[32], pk[16];
            ptx::tmem_ld_32x32b_x32(tbuf + c0, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float p0 = (c0 + j < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[j]), c, -mc)) : 0.f;
              const float p1 = (c0 + j + 1 < L) ? ptx::ex2_approx(fmaf(__uint_as_float(r[j + 1]), c, -mc)) : 0.f;
              lsum += p0 + p1;
              pk[j >> 1] = ptx::pack_bf16x2(p0, p1);
            }
            ptx::tmem_st_32x32b_x16(tbuf + (NK - ((NK - c0) >> 1)), pk);
          }
        }
        ptx::tmem_st_wait();
      }
      red_sum[wg * 128 + rit] = lsum;
      ptx::tcgen05_fence_before();
      ptx::named_bar_sync(2, 256);
      ptx::mbar_arrive(&p_ready[buf]);
      const float tot = red_sum[rit] + red_sum[128 + rit];
      // ---- epilogue: O / rowsum -> bf16 -> staging -> TMA store
      ptx::mbar_wait(&o_ready[buf], (g / C::kBufs) & 1);
      ptx::tcgen05_fence_after();
      uint32_t r[32];
      if (warp_active) {
        ptx::tmem_ld_32x32b_x32(tbuf + C::kOOff + wg * 32, r);
        ptx::tmem_ld_wait();
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(&buf_free[buf]);
      uint8_t* stage = sO + (g & 1) * C::kQBytes;
      if (warp_active) {
        const float inv = 1.0f / tot;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 u;
          u.x = ptx::pack_bf16x2(__uint_as_float(r[8 * j]) * inv, __uint_as_float(r[8 * j + 1]) * inv);
          u.y = ptx::pack_bf16x2(__uint_as_float(r[8 * j + 2]) * inv, __uint_as_float(r[8 * j + 3]) * inv);
          u.z = ptx::pack_bf16x2(__uint_as_float(r[8 * j + 4]) * inv, __uint_as_float(r[8 * j + 5]) * inv);
          u.w = ptx::pack_bf16x2(__uint_as_float(r[8 * j + 6]) * inv, __uint_as_float(r[8 * j + 7]) * inv);
          *reinterpret_cast<uint4*>(stage + rit * 128 + (((wg * 4 + j) ^ (rit & 7)) << 4)) = u;
        }
        if (wg == 0 && row < L) p.lse2[(long long)bh * L + row] = mc + log2f(tot);
      }
      ptx::fence_proxy_async();
      ptx::named_bar_sync(3, 256);
      if (issuer) {
        ptx::tma_store_3d(&tmap_o, stage, h * 64, t * 128, b);
        ptx::bulk_commit();
      }
    }
    if (issuer) ptx::bulk_wait<0>();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<C::kTmemCols>(tmem_base);
}

}  // namespace peft
