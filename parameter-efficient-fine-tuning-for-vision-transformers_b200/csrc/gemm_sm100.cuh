// Persistent warp-specialised bf16 GEMM for sm_100a:  C[M,N] = A[M,K] * B[N,K]^T  (+ fused epilogue)
//
//   * operands: bf16, K contiguous ("K-major"), fetched by TMA (cp.async.bulk.tensor, 128-byte swizzle) into a
//     kStages-deep shared-memory ring guarded by full/empty mbarriers;
//   * math: tcgen05.mma (cta_group::1, 128 x BN x 16 per instruction) issued by ONE thread, fp32 accumulators in TMEM,
//     two accumulator buffers so the epilogue of tile i overlaps the main loop of tile i+1;
//   * epilogue: 8 warps read TMEM with tcgen05.ld (one output row per thread), apply bias / GELU / QuickGELU /
//     residual / activation-derivative and store with 16-byte vector stores.
//
// This one kernel serves every dense product of the hot path (SURVEY.md §8a R1,R3,R7,R8,R10) in both directions:
// frozen weights are kept as W [N,K] (forward) and W^T [K,N] (dX), so there is never a dW product.  The adapter term
// rides in the same accumulator as extra K columns: A = [x | p], B = [W0 | scale*Up] with p = x*Dn^T the low-rank
// projection (LoRA: Dn = W1; KAdaptation: Dn[(i,c),(a,k)] = A_i[a,c] u_i[k], the reshape-trick factors), so
// W0 + dW is never materialised (R4a/R4b).
#pragma once
#include "ptx_sm100.cuh"

namespace peft {

constexpr int kBM = 128;          // rows per tile == TMEM lanes
constexpr int kBK = 64;           // bf16 per 128-byte swizzle row
constexpr int kEpiWarps = 8;      // epilogue warps 4..11
constexpr int kGemmThreads = 128 + kEpiWarps * 32;

enum GemmEpilogue : int {
  EPI_STORE = 0,       // out(bf16)   = acc (+ bias)
  EPI_GELU = 1,        // out(bf16)   = h = acc + bias ; out2(bf16) = gelu_erf(h)
  EPI_QUICKGELU = 2,   // out(bf16)   = h = acc + bias ; out2(bf16) = h * sigmoid(1.702 h)
  EPI_RESID_F32 = 3,   // out(fp32)   = aux(fp32) + acc (+ bias)
  EPI_DGELU = 4,       // out(bf16)   = acc * gelu_erf'(aux(bf16))
  EPI_DQUICKGELU = 5,  // out(bf16)   = acc * quickgelu'(aux(bf16))
  EPI_STORE_F32 = 6,   // out(fp32)   = acc (+ bias)
  EPI_RELU = 7,        // out(bf16)   = relu(acc + bias)
  EPI_DRELU = 8,       // out(bf16)   = acc * (aux(bf16) > 0)
  EPI_GELU_TANH = 9,   // out(bf16) = h ; out2 = gelu_new(h)
  EPI_DGELU_TANH = 10, // out(bf16) = acc * gelu_new'(aux)
  EPI_RESID_BF16 = 11, // out(bf16)   = aux(bf16) + acc (+ bias)
};

#ifndef PEFT_GEMM_DIAG
#define PEFT_GEMM_DIAG 0      // 1: compile the in-kernel diagnostics of the 2-CTA GEMM (make DIAG=1); see gemm2_sm100.cuh
#endif

struct GemmParams {
  int M, N, K;
  const float* bias;  // [N] fp32 or nullptr
  void* out;          // primary output
  long long ldo;      // row pitch of out, in elements
  void* out2;         // secondary output (activation) or nullptr
  long long ldo2;
  const void* aux;    // residual (fp32) / pre-activation (bf16) or nullptr
  long long ldaux;
  int ksplit = 1;     // 2-CTA kernel, EPI_STORE_F32 only: K is cut into `ksplit` ranges that are separate tiles and are
                      // summed into a pre-zeroed `out` by TMA reduce-add (weight-gradient products: small output, long K)
  // stream-K (2-CTA kernel, ksplit == 1): the first `sk_dp_tiles` tiles are processed one whole tile per cluster as usual; the
  // k-blocks of the remaining tiles are dealt out evenly over the clusters, so a tile may be computed by several clusters —
  // the one that holds its first k-block owns the epilogue and adds the others' fp32 partial accumulators, which travel
  // through `sk_ws` (one 256 x 256 fp32 slot per cluster) guarded by per-warp flags in `sk_flags`.  sk_dp_tiles < 0: off.
  int sk_dp_tiles = -1;
  float4* sk_ws = nullptr;
  int* sk_flags = nullptr;
  const float* rowscale = nullptr;   // optional fp32 [M]: (acc + bias) is multiplied by rowscale[row] before the rest of the
                      // epilogue — the per-sample stochastic-depth keep mask / keep_prob of DropPath (FS/kadaptation.py:263,276-277)
  unsigned long long* stats = nullptr;   // diagnostics (env PEFT_GEMM_STATS=1): per-CTA wait-cycle counters, see peft_debug_gemm_stats
  int dbg = 0;        // diagnostics (env PEFT_GEMM_DBG, results are WRONG when set; bench_ops.py gemm_dbg): bit 0 = no B-operand
                      // loads, bit 1 = every tile loads the operands of tile 0, bit 2 = epilogue only drains TMEM (no math / stores), bit 3 = no operand loads at all, bit 4 = no output stores, bit 5 = no activation math
};

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float dgelu_erf(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
  const float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}
__device__ __forceinline__ float quickgelu(float x) { return x / (1.0f + __expf(-1.702f * x)); }
__device__ __forceinline__ float dquickgelu(float x) {
  const float s = 1.0f / (1.0f + __expf(-1.702f * x));
  return s * (1.0f + 1.702f * x * (1.0f - s));
}
// gelu_new (transformers): 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))
__device__ __forceinline__ float gelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.0f + tanhf(u));
}
__device__ __forceinline__ float dgelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = tanhf(u);
  const float du = 0.7978845608028654f * (1.0f + 3.0f * 0.044715f * x * x);
  return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * du;
}

template <int BN>
struct GemmCfg {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
  static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N for M=128");
  static_assert(2 * BN <= 512, "two accumulators must fit TMEM");
};

template <int EPI, int NV>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, const uint32_t (&acc)[NV], long long row, int col0,
                                               bool row_ok) {
  // One thread owns NV consecutive output columns [col0, col0+NV) of one row.
  if (!row_ok) return;
  if (col0 >= p.N) return;
  const bool full = (col0 + NV <= p.N);
  float v[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) v[j] = __uint_as_float(acc[j]);
  if (p.bias != nullptr) {
    if (full) {
#pragma unroll
      for (int j = 0; j < NV; j += 4) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
        v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
      }
    } else {
      for (int j = 0; j < NV; ++j)
        if (col0 + j < p.N) v[j] += __ldg(p.bias + col0 + j);
    }
  }
  if constexpr (EPI == EPI_RESID_F32 || EPI == EPI_STORE || EPI == EPI_DGELU || EPI == EPI_DQUICKGELU)
  if (p.rowscale != nullptr) {
    const float rs = __ldg(p.rowscale + row);
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] *= rs;
  }

  if constexpr (EPI == EPI_RESID_F32 || EPI == EPI_STORE_F32) {
    float* o = reinterpret_cast<float*>(p.out) + row * p.ldo + col0;
    if constexpr (EPI == EPI_RESID_F32) {
      const float* r = reinterpret_cast<const float*>(p.aux) + row * p.ldaux + col0;
      if (full) {
#pragma unroll
        for (int j = 0; j < NV; j += 4) {
          const float4 a = *reinterpret_cast<const float4*>(r + j);
          v[j] += a.x; v[j + 1] += a.y; v[j + 2] += a.z; v[j + 3] += a.w;
        }
      } else {
        for (int j = 0; j < NV; ++j)
          if (col0 + j < p.N) v[j] += r[j];
      }
    }
    if (full) {
#pragma unroll
      for (int j = 0; j < NV; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
      for (int j = 0; j < NV; ++j)
        if (col0 + j < p.N) o[j] = v[j];
    }
    return;
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + row * p.ldo + col0;
    // aux-dependent transforms
    if constexpr (EPI == EPI_DGELU || EPI == EPI_DQUICKGELU || EPI == EPI_DRELU || EPI == EPI_DGELU_TANH ||
                  EPI == EPI_RESID_BF16) {
      const __nv_bfloat16* a = reinterpret_cast<const __nv_bfloat16*>(p.aux) + row * p.ldaux + col0;
      float h[NV];
      if (full) {
#pragma unroll
        for (int j = 0; j < NV; j += 8) {
          const uint4 u = *reinterpret_cast<const uint4*>(a + j);
          float2 f;
          f = ptx::unpack_bf16x2(u.x); h[j] = f.x; h[j + 1] = f.y;
          f = ptx::unpack_bf16x2(u.y); h[j + 2] = f.x; h[j + 3] = f.y;
          f = ptx::unpack_bf16x2(u.z); h[j + 4] = f.x; h[j + 5] = f.y;
          f = ptx::unpack_bf16x2(u.w); h[j + 6] = f.x; h[j + 7] = f.y;
        }
      } else {
        for (int j = 0; j < NV; ++j) h[j] = (col0 + j < p.N) ? __bfloat162float(a[j]) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        if constexpr (EPI == EPI_DGELU) v[j] *= dgelu_erf(h[j]);
        if constexpr (EPI == EPI_DQUICKGELU) v[j] *= dquickgelu(h[j]);
        if constexpr (EPI == EPI_DRELU) v[j] = (h[j] > 0.f) ? v[j] : 0.f;
        if constexpr (EPI == EPI_DGELU_TANH) v[j] *= dgelu_tanh(h[j]);
        if constexpr (EPI == EPI_RESID_BF16) v[j] += h[j];
      }
    }
    if constexpr (EPI == EPI_RELU) {
#pragma unroll
      for (int j = 0; j < NV; ++j) v[j] = fmaxf(v[j], 0.f);
    }
    if (full) {
#pragma unroll
      for (int j = 0; j < NV; j += 8) {
        uint4 u;
        u.x = ptx::pack_bf16x2(v[j], v[j + 1]); u.y = ptx::pack_bf16x2(v[j + 2], v[j + 3]);
        u.z = ptx::pack_bf16x2(v[j + 4], v[j + 5]); u.w = ptx::pack_bf16x2(v[j + 6], v[j + 7]);
        *reinterpret_cast<uint4*>(o + j) = u;
      }
    } else {
      for (int j = 0; j < NV; ++j)
        if (col0 + j < p.N) o[j] = __float2bfloat16(v[j]);
    }
    if constexpr (EPI == EPI_GELU || EPI == EPI_QUICKGELU || EPI == EPI_GELU_TANH) {
      __nv_bfloat16* o2 = reinterpret_cast<__nv_bfloat16*>(p.out2) + row * p.ldo2 + col0;
      // The activation is applied to the bf16-rounded pre-activation so that backward (which re-reads the stored
      // bf16 h) differentiates exactly the function the forward evaluated.
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const float hq = __bfloat162float(__float2bfloat16(v[j]));
        if constexpr (EPI == EPI_GELU) v[j] = gelu_erf(hq);
        if constexpr (EPI == EPI_QUICKGELU) v[j] = quickgelu(hq);
        if constexpr (EPI == EPI_GELU_TANH) v[j] = gelu_tanh(hq);
      }
      if (full) {
#pragma unroll
        for (int j = 0; j < NV; j += 8) {
          uint4 u;
          u.x = ptx::pack_bf16x2(v[j], v[j + 1]); u.y = ptx::pack_bf16x2(v[j + 2], v[j + 3]);
          u.z = ptx::pack_bf16x2(v[j + 4], v[j + 5]); u.w = ptx::pack_bf16x2(v[j + 6], v[j + 7]);
          *reinterpret_cast<uint4*>(o2 + j) = u;
        }
      } else {
        for (int j = 0; j < NV; ++j)
          if (col0 + j < p.N) o2[j] = __float2bfloat16(v[j]);
      }
    }
  }
}

template <int BN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte alignment.
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* tiles = smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      ptx::mbar_init(&full_bar[i], 1);
      ptx::mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tmem_full_bar[i], 1);
      ptx::mbar_init(&tmem_empty_bar[i], kEpiWarps * 32);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                        // PDL (common.cuh): no global-memory access above this line

  const int num_m = (p.M + kBM - 1) / kBM;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_kb = (p.K + kBK - 1) / kBK;

  if (warp == 0) {
    // ===================== TMA producer (converged warp; the elected lane issues) =====================
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile / num_n, n_blk = tile % num_n;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = tiles + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          if (ptx::elect_one_sync()) {
            ptx::mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
            ptx::tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kBK, m_blk * kBM);
            ptx::tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * kBK, n_blk * BN);
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (converged warp; tcgen05.mma / commit under elect.sync) =====================
    // Every lane walks the loop so that the descriptors are warp-uniform values held in uniform registers; a loop under
    // `if (lane == 0)` makes the compiler move them through an R2UR waterfall (~25 dependent instructions per MMA).
    {
      constexpr uint32_t idesc = ptx::make_idesc_bf16(kBM, BN);
      constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO 1024 B | version 1 | SWIZZLE_128B
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint32_t tiles_u32 = ptx::smem_u32(tiles);
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        ptx::mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
        ptx::tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tcgen05_fence_after();
          const uint32_t sa = tiles_u32 + stage * Cfg::kStageBytes;
          const uint32_t a_lo = ((sa & 0x3FFFFu) >> 4) | (1u << 16);
          const uint32_t b_lo = (((sa + Cfg::kABytes) & 0x3FFFFu) >> 4) | (1u << 16);
          const int krem = p.K - kb * kBK;
          const int ksteps = krem >= kBK ? (kBK / 16) : ((krem + 15) / 16);
          if (ptx::elect_one_sync()) {
            // advance 16 bf16 = 32 bytes inside the 128-byte swizzle row: +2 in the (addr >> 4) field of the low word
            ptx::umma_kblock_bf16(tmem_d, a_lo, b_lo, kDescHi, idesc, kb != 0 ? 1u : 0u, ksteps);
            ptx::umma_commit(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (ptx::elect_one_sync()) ptx::umma_commit(&tmem_full_bar[acc]);  // accumulator complete -> epilogue
        __syncwarp();
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warps =====================
    const int ew = warp - 4;
    const int quarter = warp & 3;      // TMEM lane quarter this warp may access
    const int half = ew >> 2;          // which half of the BN columns
    constexpr int kColsPerWarp = BN / 2;
    constexpr int NV = (kColsPerWarp >= 32) ? 32 : 16;
    static_assert(kColsPerWarp % NV == 0, "BN/2 must be a multiple of the chunk");
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile / num_n, n_blk = tile % num_n;
      ptx::mbar_wait(&tmem_full_bar[acc], acc_phase);
      ptx::tcgen05_fence_after();
      const long long row = static_cast<long long>(m_blk) * kBM + quarter * 32 + lane;
      const bool row_ok = row < p.M;
      const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                              static_cast<uint32_t>(acc * BN + half * kColsPerWarp);
#pragma unroll 1
      for (int c = 0; c < kColsPerWarp; c += NV) {
        uint32_t r[NV];
        if constexpr (NV == 32) ptx::tmem_ld_32x32b_x32(taddr0 + c, r);
        else ptx::tmem_ld_32x32b_x16(taddr0 + c, r);
        ptx::tmem_ld_wait();
        epilogue_chunk<EPI, NV>(p, r, row, n_blk * BN + half * kColsPerWarp + c, row_ok);
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(&tmem_empty_bar[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  pdl_trigger();                     // PDL: late on purpose (see gemm2_sm100.cuh)
  if (warp == 2) ptx::tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

}  // namespace peft
