// 2-CTA (cta_group::2) persistent bf16 GEMM for sm_100a:  C[M,N] = A[M,K] * B[N,K]^T  (+ fused epilogue)
//
// The main tensor-core kernel of the hot path (SURVEY.md §8a R3, R7, R8 and their dX products).  Compared with the
// 1-CTA kernel of gemm_sm100.cuh it
//   * pairs two SMs on one 256 x 256 output tile (tcgen05.mma.cta_group::2, M = 256): each CTA stages only its own
//     128 rows of A and 128 of the 256 B rows per k-block (32 KB instead of 48 KB), i.e. 1.5x less L2->SM operand
//     traffic per FLOP — the 1-CTA kernel measured at ~47 % of the MMA pace is bound by exactly that traffic;
//   * moves ALL epilogue HBM traffic onto TMA: accumulators go TMEM -> registers -> 128B-swizzled shared-memory slots
//     -> cp.async.bulk.tensor stores, and the epilogue's auxiliary operand (residual stream / saved pre-activation) is
//     prefetched by TMA into the same slots before the accumulator is ready.  Row-per-lane global accesses (32 sectors
//     per warp instruction) are gone; M/N tails are clipped (stores) or zero-filled (loads) by the TMA unit;
//   * evaluates the activations with MUFU-based forms (ex2 / rcp / tanh) so that the per-tile epilogue stays under
//     the per-tile MMA time at K = 768.
//
// Roles (384 threads): warp 0 = TMA producer (both CTAs), warp 1 = MMA issuer (leader CTA only), warp 2 = TMEM
// allocator, warps 4..11 = epilogue (TMEM lane quarter = warp % 4, column half = (warp - 4) / 4).
// Barriers: full/empty per smem stage (full lives in the LEADER: both CTAs' TMA loads complete on it), tmem_full /
// tmem_empty per accumulator buffer (tmem_empty in the leader collects 16 warp arrivals from both CTAs), one aux
// barrier per epilogue slot.
#pragma once
#include "gemm_sm100.cuh"

namespace peft {

constexpr int kG2Stages = 5;
constexpr int kG2BN = 256;                       // tile columns; the CTA pair computes 256 (M) x 256 (N)
constexpr int kG2StageBytes = 2 * kBM * kBK * 2; // A [128 x 64] + B half [128 x 64] bf16 = 32 KB
constexpr int kG2Threads = 128 + kEpiWarps * 32;
constexpr int kG2SlotBytes = 32 * 128;           // 32 rows x 128 B
constexpr int kG2EpiBytes = kEpiWarps * 2 * kG2SlotBytes;
constexpr int kG2BarBytes = 512;
constexpr int kG2SmemBytes = kG2Stages * kG2StageBytes + kG2EpiBytes + kG2BarBytes + 1024 /*align slack*/;

namespace ptx {
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP_C:\n"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE_C;\n"
      "bra WAIT_LOOP_C;\n"
      "WAIT_DONE_C:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// TMA load issued by either CTA of the pair; completion (bytes) is signalled on `bar_cluster_addr`, which may live in
// the peer (leader) CTA.
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* tm, uint32_t bar_cluster_addr, int c0,
                                                int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm_u32(uint32_t smem_dst, const CUtensorMap* tm, uint32_t bar_cluster_addr, int c0,
                                                    int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
// true in exactly one lane of the (fully active) warp — always the same lane, so tcgen05.commit tracks the MMAs it issued
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "selp.b32 %0, 1, 0, P;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
// One 64-wide k-block of the 2-CTA product: up to four tcgen05.mma (K = 16 each) whose shared-memory descriptors differ only
// in the start-address field of the low word (+step per K-step).  The first MMA accumulates iff acc_first != 0.
__device__ __forceinline__ void umma_kblock_bf16_2sm(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                                     uint32_t acc_first, int ksteps, uint32_t step) {
  asm volatile(
      "{\n"
      ".reg .pred p0, p1, p2, p3, pt;\n"
      ".reg .b64 da, db;\n"
      ".reg .b32 al, bl;\n"
      "setp.ne.b32 p0, %5, 0;\n"
      "setp.gt.s32 p1, %6, 1;\n"
      "setp.gt.s32 p2, %6, 2;\n"
      "setp.gt.s32 p3, %6, 3;\n"
      "setp.eq.u32 pt, %7, %7;\n"
      "mov.b64 da, {%1, %3};\n"
      "mov.b64 db, {%2, %3};\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p0;\n"
      "add.u32 al, %1, %7;\n"
      "add.u32 bl, %2, %7;\n"
      "mov.b64 da, {al, %3};\n"
      "mov.b64 db, {bl, %3};\n"
      "@p1 tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, pt;\n"
      "add.u32 al, al, %7;\n"
      "add.u32 bl, bl, %7;\n"
      "mov.b64 da, {al, %3};\n"
      "mov.b64 db, {bl, %3};\n"
      "@p2 tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, pt;\n"
      "add.u32 al, al, %7;\n"
      "add.u32 bl, bl, %7;\n"
      "mov.b64 da, {al, %3};\n"
      "mov.b64 db, {bl, %3};\n"
      "@p3 tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, pt;\n"
      "}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(acc_first), "r"(ksteps), "r"(step)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
// MN-major SWIZZLE_128B operand, 128 MN-elements wide: two 64-wide atoms columns 8 KB apart (LBO), 8-row groups 1 KB apart (SBO)
__device__ __forceinline__ uint64_t make_smem_desc_sw128_mn128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(8192 >> 4) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
// same box, but added into global memory (fp32 tensor map): split-K partial sums
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* tm, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_slot) {  // one whole warp in EACH CTA of the pair
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on the mbarrier at this shared-memory offset in every CTA of `mask` once all prior MMAs have completed.
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
}  // namespace ptx

// ---- MUFU-based activations.  Phi(x) = sigmoid(2 x q(x^2)) with q fitted to atanh(erf(x / sqrt 2)) / x:
// max |x Phi_fit(x) - gelu_erf(x)| = 2.6e-5 over the reals (a bf16 ulp at |gelu| ~ 1 is 3.9e-3).
__device__ __forceinline__ float phi_u(float x) {  // returns u = x q(x^2); Phi = sigmoid(2u) = 0.5 + 0.5 tanh(u)
  const float s = fminf(x * x, 50.0f);   // q peaks at s = 52; beyond |x| = 7 the sigmoid is saturated anyway
  return x * fmaf(s, fmaf(s, -3.53110710e-4f, 3.70155061e-2f), 7.97496957e-1f);
}
__device__ __forceinline__ float fast_gelu_erf(float x) {
  const float e = ptx::ex2_approx(-2.8853900817779268f * phi_u(x));   // exp(-2u)
  return x * ptx::rcp_approx(1.0f + e);
}
__device__ __forceinline__ float fast_dgelu_erf(float x) {
  const float cdf = fmaf(0.5f, ptx::tanh_approx(phi_u(x)), 0.5f);
  const float pdf = 0.3989422804014327f * ptx::ex2_approx(-0.72134752044448170f * x * x);
  return fmaf(x, pdf, cdf);
}
__device__ __forceinline__ float fast_quickgelu(float x) {
  return x * ptx::rcp_approx(1.0f + ptx::ex2_approx(-2.4554669595930156f * x));   // 1.702 * log2(e)
}
__device__ __forceinline__ float fast_dquickgelu(float x) {
  const float s = ptx::rcp_approx(1.0f + ptx::ex2_approx(-2.4554669595930156f * x));
  return s * fmaf(1.702f * x, 1.0f - s, 1.0f);
}
__device__ __forceinline__ float fast_gelu_tanh(float x) {
  const float u = 0.7978845608028654f * fmaf(0.044715f * x * x, x, x);
  return x * ptx::rcp_approx(1.0f + ptx::ex2_approx(-2.8853900817779268f * u));
}
__device__ __forceinline__ float fast_dgelu_tanh(float x) {
  const float x2 = x * x;
  const float u = 0.7978845608028654f * fmaf(0.044715f * x2, x, x);
  const float t = ptx::tanh_approx(u);
  const float du = 0.7978845608028654f * fmaf(0.134145f, x2, 1.0f);
  return fmaf(0.5f * x * (1.0f - t * t), du, fmaf(0.5f, t, 0.5f));
}

// ---- packed fp32 arithmetic (sm_100 FFMA2 / FMUL2 / FADD2: two fp32 lanes per instruction, full precision).  The epilogues
// of the activation products are issue-bound with two warps per scheduler, so halving the FP instruction count is what
// brings the per-tile epilogue under the per-tile MMA time (r02: fc1+GELU 18 -> ~9 instructions per element).
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("{\n.reg .b64 ra, rb, rc, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nmov.b64 rc, {%6, %7};\n"
      "fma.rn.f32x2 rd, ra, rb, rc;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return d;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  float2 d;
  asm("{\n.reg .b64 ra, rb, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nmul.rn.f32x2 rd, ra, rb;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  float2 d;
  asm("{\n.reg .b64 ra, rb, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nadd.rn.f32x2 rd, ra, rb;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 splat2(float a) { return make_float2(a, a); }
__device__ __forceinline__ float2 tanh2(float2 u) { return make_float2(ptx::tanh_approx(u.x), ptx::tanh_approx(u.y)); }

// gelu_erf(x) ~ x (0.5 + 0.5 tanh(u)), u = x q(min(x^2, 50)) — the same fit as phi_u, one MUFU per element
__device__ __forceinline__ float2 gelu_u2(float2 x, float2& s) {
  s = mul2(x, x);
  s.x = fminf(s.x, 50.0f);
  s.y = fminf(s.y, 50.0f);
  const float2 q = fma2(s, fma2(s, splat2(-3.53110710e-4f), splat2(3.70155061e-2f)), splat2(7.97496957e-1f));
  return mul2(x, q);
}
__device__ __forceinline__ float2 gelu_erf2(float2 x) {
  float2 s;
  const float2 t = tanh2(gelu_u2(x, s));
  const float2 hx = mul2(x, splat2(0.5f));
  return fma2(hx, t, hx);
}
// exact derivative of the fitted function above (so backward differentiates what forward evaluated):
// d/dx = 0.5 + 0.5 t + 0.5 x (1 - t^2) u'(x),  u' = q + 2 s q'(s) = c0 + 3 c1 s + 5 c2 s^2  (in the clamped range 1 - t^2 = 0)
__device__ __forceinline__ float2 dgelu_erf2(float2 x) {
  float2 s;
  const float2 t = tanh2(gelu_u2(x, s));
  const float2 du = fma2(s, fma2(s, splat2(5.0f * -3.53110710e-4f), splat2(3.0f * 3.70155061e-2f)), splat2(7.97496957e-1f));
  const float2 omt = fma2(mul2(t, splat2(-1.0f)), t, splat2(1.0f));
  const float2 z = mul2(mul2(x, splat2(0.5f)), du);
  return fma2(z, omt, fma2(t, splat2(0.5f), splat2(0.5f)));
}
// x sigmoid(1.702 x) = x (0.5 + 0.5 tanh(0.851 x))
__device__ __forceinline__ float2 quickgelu2(float2 x) {
  const float2 t = tanh2(mul2(x, splat2(0.851f)));
  const float2 hx = mul2(x, splat2(0.5f));
  return fma2(hx, t, hx);
}
// sigma + 1.702 x sigma (1 - sigma), sigma = 0.5 + 0.5 t:  sigma (1 - sigma) = 0.25 (1 - t^2)
__device__ __forceinline__ float2 dquickgelu2(float2 x) {
  const float2 t = tanh2(mul2(x, splat2(0.851f)));
  const float2 omt = fma2(mul2(t, splat2(-1.0f)), t, splat2(1.0f));
  return fma2(mul2(x, splat2(0.25f * 1.702f)), omt, fma2(t, splat2(0.5f), splat2(0.5f)));
}
// gelu_new: 0.5 x (1 + tanh(x (0.79788 + 0.035677 x^2)))
__device__ __forceinline__ float2 gelu_tanh2(float2 x) {
  const float2 s = mul2(x, x);
  const float2 t = tanh2(mul2(x, fma2(s, splat2(0.7978845608028654f * 0.044715f), splat2(0.7978845608028654f))));
  const float2 hx = mul2(x, splat2(0.5f));
  return fma2(hx, t, hx);
}
__device__ __forceinline__ float2 dgelu_tanh2(float2 x) {
  const float2 s = mul2(x, x);
  const float2 t = tanh2(mul2(x, fma2(s, splat2(0.7978845608028654f * 0.044715f), splat2(0.7978845608028654f))));
  const float2 du = fma2(s, splat2(3.0f * 0.7978845608028654f * 0.044715f), splat2(0.7978845608028654f));
  const float2 omt = fma2(mul2(t, splat2(-1.0f)), t, splat2(1.0f));
  return fma2(mul2(mul2(x, splat2(0.5f)), du), omt, fma2(t, splat2(0.5f), splat2(0.5f)));
}

template <int EPI>
struct EpiTraits {
  static constexpr bool kF32Out = (EPI == EPI_RESID_F32 || EPI == EPI_STORE_F32);
  static constexpr bool kAux = (EPI == EPI_RESID_F32 || EPI == EPI_DGELU || EPI == EPI_DQUICKGELU || EPI == EPI_DRELU ||
                                EPI == EPI_DGELU_TANH || EPI == EPI_RESID_BF16);
  static constexpr bool kDual = (EPI == EPI_GELU || EPI == EPI_QUICKGELU || EPI == EPI_GELU_TANH);
  static constexpr int kUnitCols = kF32Out ? 32 : 64;    // columns per 128-byte slot row
  static constexpr int kUnits = 128 / kUnitCols;         // units per warp per tile (the warp owns 128 columns)
};

// Tile scheduler shared by the three roles of the kernel.  Work items ("segments") are (tile, [kb0, kb1), kind):
//   kSegFull     a whole tile, or — with GemmParams::ksplit > 1 — one K range of a tile that is reduce-added into `out`;
//   kSegOwner    stream-K: the first k-blocks of a tile; the epilogue adds the partial sums of the following cluster(s);
//   kSegPartial  stream-K: later k-blocks of a tile; the accumulator is handed to the owner through the workspace.
// A cluster's stream-K range starts with at most one kSegPartial and ends with at most one kSegOwner, so a waiting owner
// only ever depends on segments that other clusters process FIRST in their range: no cyclic waits.
enum : int { kSegFull = 0, kSegOwner = 1, kSegPartial = 2 };
struct TileSched {
  int cid, C, ksplit, num_kb, kb_per, n_dp, i;
  long long u, u1;        // stream-K cursor / end, in k-block units over the stream-K tiles
  int sk_base;            // first stream-K tile
  __device__ TileSched(const GemmParams& p, int num_tiles_mn, int num_kb_, int cluster_id, int num_clusters) {
    cid = cluster_id; C = num_clusters; num_kb = num_kb_; i = 0;
    ksplit = p.ksplit > 1 ? p.ksplit : 1;
    kb_per = (num_kb + ksplit - 1) / ksplit;
    if (p.sk_dp_tiles >= 0 && ksplit == 1) {
      n_dp = p.sk_dp_tiles;
      sk_base = p.sk_dp_tiles;
      const long long sku = (long long)(num_tiles_mn - p.sk_dp_tiles) * num_kb;
      u = sku * cid / C;
      u1 = sku * (cid + 1) / C;
    } else {
      n_dp = num_tiles_mn * ksplit;
      sk_base = 0;
      u = u1 = 0;
    }
  }
  // first unit of cluster c's stream-K range (for the owner's contributor walk)
  __device__ long long range_begin(const GemmParams& p, int num_tiles_mn, int c) const {
    return (long long)(num_tiles_mn - p.sk_dp_tiles) * num_kb * c / C;
  }
  __device__ bool next(int& mn, int& kb0, int& kb1, int& kind) {
    const int t = cid + i * C;
    if (t < n_dp) {
      ++i;
      mn = t / ksplit;
      kb0 = (t % ksplit) * kb_per;
      kb1 = min(num_kb, kb0 + kb_per);
      kind = kSegFull;
      return true;
    }
    if (u >= u1) return false;
    const int st = (int)(u / num_kb);
    kb0 = (int)(u - (long long)st * num_kb);
    const long long left = u1 - u;
    kb1 = (left >= (long long)(num_kb - kb0)) ? num_kb : kb0 + (int)left;
    mn = sk_base + st;
    kind = (kb0 == 0) ? (kb1 == num_kb ? kSegFull : kSegOwner) : kSegPartial;
    u += kb1 - kb0;
    return true;
  }
};

// 16-byte chunk j of row `lane` inside a 128B-swizzled [32 x 128 B] slot (the layout TMA expects for SWIZZLE_128B).
__device__ __forceinline__ uint32_t slot_off(int lane, int j) { return (uint32_t)(lane * 128 + ((j ^ (lane & 7)) << 4)); }

// TN = true: both operands are given token-major — A(m, k) = a[k, m], B(n, k) = b[k, n] (the weight-gradient product
// dW = dY^T X straight from the row-major activations, no transposes).  The smem tiles are then MN-major: a TMA box is
// [64 k-rows x 64 MN-elements (128 B)], two boxes side by side give the 128-wide operand half of a CTA (second box at
// +8 KB = the descriptor's leading-dimension byte offset), 8-row groups stay 1 KB apart, and one UMMA K-step (16 rows)
// advances the start address by 2 KB.  p.K is the number of tokens; rows past it are zero-filled by TMA.
template <int EPI, bool TN = false>
__global__ void __launch_bounds__(kG2Threads, 1)
gemm2_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                          const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_out2,
                          const __grid_constant__ CUtensorMap tmap_aux, const GemmParams p) {
  using T = EpiTraits<EPI>;
  // Diagnostics (wait-cycle / phase counters, PEFT_GEMM_DBG modes) are compiled in only with -DPEFT_GEMM_DIAG=1 (make DIAG=1): in
  // the production build the conditions below are compile-time false and the code disappears — a CTA of the fixed-global-batch
  // split runs one to four tiles, so instruction-cache-cold code is part of its launch floor.
#if PEFT_GEMM_DIAG
  unsigned long long* const dstats = p.stats;
  const int ddbg = p.dbg;
#else
  constexpr unsigned long long* dstats = nullptr;
  constexpr int ddbg = 0;
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* tiles = smem;
  uint8_t* epi_smem = smem + kG2Stages * kG2StageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_smem + kG2EpiBytes);
  uint64_t* empty_bar = full_bar + kG2Stages;
  uint64_t* tmem_full_bar = empty_bar + kG2Stages;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;      // [2]
  uint64_t* aux_bar = tmem_empty_bar + 2;            // [kEpiWarps][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux_bar + 2 * kEpiWarps);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = ptx::cluster_ctarank();
  const bool leader = cta_rank == 0;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
    ptx::prefetch_tmap(&tmap_out);
    if constexpr (T::kDual) ptx::prefetch_tmap(&tmap_out2);
    if constexpr (T::kAux) ptx::prefetch_tmap(&tmap_aux);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kG2Stages; ++i) {
      ptx::mbar_init(&full_bar[i], 1);    // leader producer's arrive.expect_tx (used in the leader only)
      ptx::mbar_init(&empty_bar[i], 1);   // one multicast tcgen05.commit
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tmem_full_bar[i], 1);
      ptx::mbar_init(&tmem_empty_bar[i], 2 * kEpiWarps);   // every epilogue warp of both CTAs (leader's copy is used)
    }
    for (int i = 0; i < 2 * kEpiWarps; ++i) ptx::mbar_init(&aux_bar[i], 1);
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc_2sm<512>(tmem_slot);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::cluster_sync();               // peer barriers initialised + peer TMEM allocated before any cross-CTA traffic
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                        // everything above overlapped the predecessor's tail; global memory only from here on

  const int num_mp = (p.M + 2 * kBM - 1) / (2 * kBM);     // 256-row pair tiles
  const int num_n = (p.N + kG2BN - 1) / kG2BN;
  const int ksplit = p.ksplit > 1 ? p.ksplit : 1;
  const int num_tiles_mn = num_mp * num_n;                  // output tiles; the scheduler adds K splits / stream-K segments
  const int num_kb = (p.K + kBK - 1) / kBK;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  // Warps 0 and 1 run their loops CONVERGED (all 32 lanes) and only the TMA / MMA / commit instructions sit under an
  // elect.sync predicate: every operand of those instructions is then provably warp-uniform and lives in uniform registers.
  // With the whole loop inside `if (lane == 0)` the compiler has to move each descriptor into uniform registers through an
  // ELECT / R2UR.BROADCAST waterfall loop (~25 dependent scalar instructions per tcgen05.mma), and that chain — not the
  // tensor pipe, the operand feed or the epilogue — paced the kernel at ~200 clk per 128-clk UMMA (r02 gemm_probe).
  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    int stage = 0;
    uint32_t phase = 0;
    long long t_wait = 0, t_begin = clock64();
    const uint32_t tiles_u32 = ptx::smem_u32(tiles);
    const uint32_t full0 = ptx::mapa(ptx::smem_u32(&full_bar[0]), 0);     // the leader's full barriers (8 bytes apart)
    TileSched sched(p, num_tiles_mn, num_kb, cluster_id, num_clusters);
    int mn, kb0, kb1, kind;
    while (sched.next(mn, kb0, kb1, kind)) {
      const int mp = mn / num_n, n_blk = mn % num_n;
      const int row_a = (ddbg & 2) ? (int)cta_rank * kBM : mp * 2 * kBM + (int)cta_rank * kBM;
      const int row_b = (ddbg & 2) ? (int)cta_rank * (kG2BN / 2) : n_blk * kG2BN + (int)cta_rank * (kG2BN / 2);
      for (int kb = kb0; kb < kb1; ++kb) {
        if (dstats) {
          const long long t0 = clock64();
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          t_wait += clock64() - t0;
        } else {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        }
        const uint32_t sa = tiles_u32 + stage * kG2StageBytes;
        const uint32_t sb = sa + kBM * kBK * 2;
        const uint32_t bar = full0 + stage * 8;
        if (ptx::elect_one()) {
          if (ddbg & 8) {                   // diagnostic: no operand loads at all (pure MMA + epilogue on stale smem)
            if (leader) ptx::mbar_expect_tx(&full_bar[stage], 0);
          } else if (!TN && (ddbg & 1)) {   // diagnostic: A operand only (the B slot keeps stale data)
            if (leader) ptx::mbar_expect_tx(&full_bar[stage], kG2StageBytes);
            ptx::tma_load_2d_2sm_u32(sa, &tmap_a, bar, kb * kBK, row_a);
          } else {
            if (leader) ptx::mbar_expect_tx(&full_bar[stage], 2 * kG2StageBytes);
            if constexpr (TN) {
              ptx::tma_load_2d_2sm_u32(sa, &tmap_a, bar, row_a, kb * kBK);
              ptx::tma_load_2d_2sm_u32(sa + 8192, &tmap_a, bar, row_a + 64, kb * kBK);
              ptx::tma_load_2d_2sm_u32(sb, &tmap_b, bar, row_b, kb * kBK);
              ptx::tma_load_2d_2sm_u32(sb + 8192, &tmap_b, bar, row_b + 64, kb * kBK);
            } else {
              ptx::tma_load_2d_2sm_u32(sa, &tmap_a, bar, kb * kBK, row_a);
              ptx::tma_load_2d_2sm_u32(sb, &tmap_b, bar, kb * kBK, row_b);
            }
          }
        }
        __syncwarp();
        if (++stage == kG2Stages) { stage = 0; phase ^= 1; }
      }
    }
    if (dstats && !(ddbg & 128) && lane == 0) { dstats[blockIdx.x * 8 + 0] = t_wait; dstats[blockIdx.x * 8 + 1] = clock64() - t_begin; }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA; one elected lane issues) =====================
    if (leader) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16(2 * kBM, kG2BN) | (TN ? ((1u << 15) | (1u << 16)) : 0u);
      // shared-memory matrix descriptor = {lo: start address >> 4 | LBO field, hi: SBO 1024 B | version 1 | SWIZZLE_128B}; the
      // K-step advance (32 B inside the swizzle row, or 2 KB for the MN-major TN tiles) only touches the low word
      constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);
      constexpr uint32_t kDescLoFlags = TN ? ((8192u >> 4) << 16) : (1u << 16);
      constexpr uint32_t kStepUnits = TN ? (2048u >> 4) : (32u >> 4);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      long long t_wfull = 0, t_wtmem = 0, t_begin = clock64();
      const uint32_t tiles_u32 = ptx::smem_u32(tiles);
      TileSched sched(p, num_tiles_mn, num_kb, cluster_id, num_clusters);
      int mn, kb0, kb1, kind;
      while (sched.next(mn, kb0, kb1, kind)) {
        if (dstats) {
          const long long t0 = clock64();
          ptx::mbar_wait_cluster(&tmem_empty_bar[acc], acc_phase ^ 1);
          t_wtmem += clock64() - t0;
        } else {
          ptx::mbar_wait_cluster(&tmem_empty_bar[acc], acc_phase ^ 1);
        }
        ptx::tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * kG2BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          if (dstats) {
            const long long t0 = clock64();
            ptx::mbar_wait_cluster(&full_bar[stage], phase);
            t_wfull += clock64() - t0;
          } else {
            ptx::mbar_wait_cluster(&full_bar[stage], phase);
          }
          ptx::tcgen05_fence_after();
          const uint32_t sa = tiles_u32 + stage * kG2StageBytes;
          const uint32_t a_lo = ((sa & 0x3FFFFu) >> 4) | kDescLoFlags;
          const uint32_t b_lo = (((sa + kBM * kBK * 2) & 0x3FFFFu) >> 4) | kDescLoFlags;
          const int krem = p.K - kb * kBK;
          const int ksteps = krem >= kBK ? (kBK / 16) : ((krem + 15) / 16);
          if (ptx::elect_one()) {
            ptx::umma_kblock_bf16_2sm(tmem_d, a_lo, b_lo, kDescHi, idesc, kb != kb0 ? 1u : 0u, ksteps, kStepUnits);
            ptx::umma_commit_2sm(&empty_bar[stage], 3);       // frees this smem slot in both CTAs
          }
          __syncwarp();
          if (++stage == kG2Stages) { stage = 0; phase ^= 1; }
        }
        if (ptx::elect_one()) ptx::umma_commit_2sm(&tmem_full_bar[acc], 3);   // accumulator complete -> both CTAs' epilogues
        __syncwarp();
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      if (dstats && !(ddbg & 128) && lane == 0) {
        dstats[blockIdx.x * 8 + 2] = t_wfull; dstats[blockIdx.x * 8 + 3] = t_wtmem; dstats[blockIdx.x * 8 + 4] = clock64() - t_begin;
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warps =====================
    const int ew = warp - 4;
    const int quarter = warp & 3;
    const int half = ew >> 2;
    uint8_t* const slot_base = epi_smem + ew * 2 * kG2SlotBytes;
    auto slot_ptr = [&](int s) { return slot_base + s * kG2SlotBytes; };
    uint64_t* abar = aux_bar + 2 * ew;
    uint32_t aux_phase = 0;   // bit s = parity of slot s
    const uint32_t tmem_empty_remote0 = ptx::mapa(ptx::smem_u32(&tmem_empty_bar[0]), 0);
    const uint32_t tmem_empty_remote1 = ptx::mapa(ptx::smem_u32(&tmem_empty_bar[1]), 0);
    int acc = 0;
    uint32_t acc_phase = 0;
    long long t_wacc = 0, t_begin = clock64();
    // diagnostics (PEFT_GEMM_STATS + PEFT_GEMM_DBG bit 7): cycles of epilogue warp 0 per phase — 0 TMEM load, 1 bias / row scale,
    // 2 slot or aux wait (+ aux math), 3 pack + st.shared + fence + store issue, 4 activation math
    const bool ph = dstats != nullptr && (ddbg & 128);
    long long tp[5] = {0, 0, 0, 0, 0}, tc = 0;
    auto tick = [&](int i) { if (ph) { const long long n = clock64(); tp[i] += n - tc; tc = n; } };
    TileSched sched(p, num_tiles_mn, num_kb, cluster_id, num_clusters);
    // stream-K workspace of this warp inside cluster c's slot: [2 CTAs][8 warps][1024 float4], unit-major then 16-byte chunk
    // then lane (coalesced); sk_flags[(c * 2 + cta) * 8 + warp] = 1 once that warp's part of the partial tile is visible
    auto sk_slot = [&](int c) { return p.sk_ws + ((size_t)(c * 2 + (int)cta_rank) * kEpiWarps + ew) * 1024; };
    auto sk_flag = [&](int c) { return p.sk_flags + (c * 2 + (int)cta_rank) * kEpiWarps + ew; };
    int mn, kb0, kb1, kind;
    while (sched.next(mn, kb0, kb1, kind)) {
      const int mp = mn / num_n, n_blk = mn % num_n;
      const int row0 = mp * 2 * kBM + (int)cta_rank * kBM + quarter * 32;   // first row of this warp's 32-row slab
      const int col0 = n_blk * kG2BN + half * 128;                          // first of this warp's 128 columns
      // stream-K owner: clusters cluster_id + 1 .. sk_last hold the rest of this tile's K range
      int sk_last = cluster_id;
      if (kind == kSegOwner) {
        const long long tile_end = (long long)(mn - p.sk_dp_tiles + 1) * num_kb;
        while (sk_last + 1 < num_clusters && sched.range_begin(p, num_tiles_mn, sk_last + 1) < tile_end) ++sk_last;
      }
      if (T::kAux && kind != kSegPartial) {
        // prefetch the auxiliary operand of units 0 and 1 (independent of the accumulator)
        if (lane == 0) {
          ptx::bulk_wait_read<0>();                   // previous tile's stores have released both slots
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            ptx::mbar_expect_tx(&abar[s], kG2SlotBytes);
            ptx::tma_load_2d(slot_ptr(s), &tmap_aux, &abar[s], col0 + s * T::kUnitCols, row0);
          }
        }
      }
      if (dstats) {
        const long long t0 = clock64();
        ptx::mbar_wait(&tmem_full_bar[acc], acc_phase);
        t_wacc += clock64() - t0;
      } else {
        ptx::mbar_wait(&tmem_full_bar[acc], acc_phase);
      }
      ptx::tcgen05_fence_after();
      const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                              static_cast<uint32_t>(acc * kG2BN + half * 128);
      if (kind == kSegOwner) {
        // the contributors' parts of this tile (same CTA rank, same warp role) must be visible before they are added below
        if (lane == 0) {
          for (int c = cluster_id + 1; c <= sk_last; ++c) {
            unsigned spins = 0;
            int f;
            do {
              asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(f) : "l"(sk_flag(c)) : "memory");
              if (f == 0 && ++spins > (1u << 28)) __trap();     // watchdog: a lost partial must not hang the GPU
            } while (f == 0);
          }
        }
        __syncwarp();
      }
#pragma unroll 1
      for (int u = 0; u < T::kUnits; ++u) {
        const int s = u & 1;
        const int colu = col0 + u * T::kUnitCols;
        constexpr int NP = T::kUnitCols / 2;     // the unit as fp32 PAIRS: all arithmetic below is packed (FFMA2 / FMUL2 / FADD2)
        float2 v[NP];
        if (ph) tc = clock64();
        {
          uint32_t r[32];
          ptx::tmem_ld_32x32b_x32(taddr0 + u * T::kUnitCols, r);
          if constexpr (T::kUnitCols == 64) {
            uint32_t r2[32];
            ptx::tmem_ld_32x32b_x32(taddr0 + u * T::kUnitCols + 32, r2);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              v[j] = make_float2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
              v[16 + j] = make_float2(__uint_as_float(r2[2 * j]), __uint_as_float(r2[2 * j + 1]));
            }
          } else {
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = make_float2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
          }
        }
        tick(0);
        if (u == T::kUnits - 1) {
          // all of this warp's accumulator reads are done: hand the TMEM buffer back to the MMA issuer
          ptx::tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive_cluster(acc ? tmem_empty_remote1 : tmem_empty_remote0);
        }
        if constexpr (!T::kAux) { if (ddbg & 4) continue; }   // diagnostic: drain TMEM only
        if (kind == kSegPartial) {
          // stream-K: raw fp32 accumulator -> this cluster's workspace slot (L2-resident, read back once by the owner)
          float4* ws = sk_slot(cluster_id) + u * (NP / 2) * 32 + lane;
#pragma unroll
          for (int jj = 0; jj < NP / 2; ++jj) __stcg(ws + jj * 32, make_float4(v[2 * jj].x, v[2 * jj].y, v[2 * jj + 1].x, v[2 * jj + 1].y));
          continue;
        }
        if (kind == kSegOwner) {
          for (int c = cluster_id + 1; c <= sk_last; ++c) {
            const float4* ws = sk_slot(c) + u * (NP / 2) * 32 + lane;
#pragma unroll
            for (int jj = 0; jj < NP / 2; ++jj) {
              const float4 a = __ldcg(ws + jj * 32);
              v[2 * jj] = add2(v[2 * jj], make_float2(a.x, a.y));
              v[2 * jj + 1] = add2(v[2 * jj + 1], make_float2(a.z, a.w));
            }
          }
        }
        if (p.bias != nullptr) {
          if (colu + T::kUnitCols <= p.N) {
#pragma unroll
            for (int j = 0; j < NP; j += 2) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + colu + 2 * j));
              v[j] = add2(v[j], make_float2(b.x, b.y));
              v[j + 1] = add2(v[j + 1], make_float2(b.z, b.w));
            }
          } else {
#pragma unroll
            for (int j = 0; j < NP; ++j) {
              if (colu + 2 * j < p.N) v[j].x += __ldg(p.bias + colu + 2 * j);
              if (colu + 2 * j + 1 < p.N) v[j].y += __ldg(p.bias + colu + 2 * j + 1);
            }
          }
        }
        // stochastic-depth row scale, lane = row (re-read per unit: no register held across units); only the epilogues the
        // engine uses it with are compiled with it (the host rejects the others)
        if constexpr (EPI == EPI_RESID_F32 || EPI == EPI_STORE || EPI == EPI_DGELU || EPI == EPI_DQUICKGELU)
        if (p.rowscale != nullptr) {
          const float2 rs = splat2((row0 + lane < p.M) ? __ldg(p.rowscale + row0 + lane) : 0.0f);
#pragma unroll
          for (int j = 0; j < NP; ++j) v[j] = mul2(v[j], rs);
        }
        tick(1);
        // ---- slot availability / auxiliary operand
        if constexpr (T::kAux) {
          ptx::mbar_wait(&abar[s], (aux_phase >> s) & 1u);
          aux_phase ^= (1u << s);
          if constexpr (EPI == EPI_RESID_F32) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 a = *reinterpret_cast<const float4*>(slot_ptr(s) + slot_off(lane, j));
              v[2 * j] = add2(v[2 * j], make_float2(a.x, a.y));
              v[2 * j + 1] = add2(v[2 * j + 1], make_float2(a.z, a.w));
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint4 a = *reinterpret_cast<const uint4*>(slot_ptr(s) + slot_off(lane, j));
              const uint32_t au[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float2 h = ptx::unpack_bf16x2(au[q]);
                float2& vv = v[4 * j + q];
                if constexpr (EPI == EPI_DGELU) vv = mul2(vv, dgelu_erf2(h));
                if constexpr (EPI == EPI_DQUICKGELU) vv = mul2(vv, dquickgelu2(h));
                if constexpr (EPI == EPI_DGELU_TANH) vv = mul2(vv, dgelu_tanh2(h));
                if constexpr (EPI == EPI_DRELU) { vv.x = h.x > 0.f ? vv.x : 0.f; vv.y = h.y > 0.f ? vv.y : 0.f; }
                if constexpr (EPI == EPI_RESID_BF16) vv = add2(vv, h);
              }
            }
          }
          __syncwarp();   // every lane has read its aux row before anyone overwrites the slot
        } else if constexpr (T::kDual) {
          if (lane == 0) ptx::bulk_wait_read<1>();   // slot 0's previous store (h of the previous unit) has been read; the
          __syncwarp();                              // activation store issued after it may still be draining slot 1
        } else {
          if (lane == 0) ptx::bulk_wait_read<1>();   // the store issued two units ago (same slot) has been read
          __syncwarp();
        }
        if constexpr (EPI == EPI_RELU) {
#pragma unroll
          for (int j = 0; j < NP; ++j) { v[j].x = fmaxf(v[j].x, 0.f); v[j].y = fmaxf(v[j].y, 0.f); }
        }
        tick(2);
        // ---- registers -> swizzled slot
        if constexpr (T::kF32Out) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(slot_ptr(s) + slot_off(lane, j)) = make_float4(v[2 * j].x, v[2 * j].y, v[2 * j + 1].x, v[2 * j + 1].y);
        } else {
          uint8_t* dst = T::kDual ? slot_ptr(0) : slot_ptr(s);
          uint32_t pk[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) pk[j] = ptx::pack_bf16x2(v[j].x, v[j].y);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(dst + slot_off(lane, j)) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
          if constexpr (T::kDual) {
            // the pre-activation store leaves now, so that its shared-memory read drains under the activation math
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0 && !(ddbg & 16)) {
              ptx::tma_store_2d(&tmap_out, slot_ptr(0), colu, row0);
              ptx::bulk_commit();
            }
            tick(3);
            // the activation is evaluated on the fp32 pre-activation (its bf16 copy goes to `out` for the backward pass: the
            // two differ by one bf16 rounding of h, 2^-9 relative, far inside the activation's own bf16 output rounding)
            // (the diagnostic switch sits OUTSIDE the unrolled loop: a branch per pair splits it into 32 basic blocks, and each
            // pair's mul -> fma -> fma -> MUFU -> fma -> pack chain then runs at full latency with nothing to overlap — measured
            // 77 clk per pair per warp, 2.5 k clk per 64 columns, r02 gemm_phases)
            if (ddbg & 32) {                                                              // diagnostic: no activation math
#pragma unroll
              for (int j = 0; j < 32; ++j) pk[j] = ptx::pack_bf16x2(v[j].y, v[j].x);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                float2 a;
                if constexpr (EPI == EPI_GELU) a = gelu_erf2(v[j]);
                else if constexpr (EPI == EPI_QUICKGELU) a = quickgelu2(v[j]);
                else a = gelu_tanh2(v[j]);
                pk[j] = ptx::pack_bf16x2(a.x, a.y);
              }
            }
            tick(4);
            if (lane == 0) ptx::bulk_wait_read<1>();   // slot 1's previous store has been read (only the h store above is younger)
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *reinterpret_cast<uint4*>(slot_ptr(1) + slot_off(lane, j)) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
          }
        }
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (ddbg & 16) {                          // diagnostic: no output stores
          } else if constexpr (T::kDual) {
            ptx::tma_store_2d(&tmap_out2, slot_ptr(1), colu, row0);
          } else if (EPI == EPI_STORE_F32 && ksplit > 1) {
            ptx::tma_reduce_add_2d(&tmap_out, slot_ptr(s), colu, row0);
          } else {
            ptx::tma_store_2d(&tmap_out, slot_ptr(s), colu, row0);
          }
          ptx::bulk_commit();
          if constexpr (T::kAux) {
            if (u + 2 < T::kUnits) {
              ptx::bulk_wait_read<0>();               // the store just issued has drained the slot
              ptx::mbar_expect_tx(&abar[s], kG2SlotBytes);
              ptx::tma_load_2d(slot_ptr(s), &tmap_aux, &abar[s], col0 + (u + 2) * T::kUnitCols, row0);
            }
          }
        }
        __syncwarp();
        tick(3);
      }
      if (kind == kSegPartial) {
        __syncwarp();                         // every lane's partial-sum stores are ordered before lane 0's fence + flag
        if (lane == 0) {
          __threadfence();
          asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(sk_flag(cluster_id)), "r"(1) : "memory");
        }
      } else if (kind == kSegOwner) {
        __syncwarp();                         // all lanes have read the partials: re-arm the flags for the next launch
        if (lane == 0)
          for (int c = cluster_id + 1; c <= sk_last; ++c) *reinterpret_cast<volatile int*>(sk_flag(c)) = 0;
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (lane == 0) ptx::bulk_wait<0>();   // all global writes of this warp are complete before the CTA exits
    if (dstats && ew == 0 && lane == 0) {
      dstats[blockIdx.x * 8 + 5] = t_wacc; dstats[blockIdx.x * 8 + 6] = clock64() - t_begin;
      if (ph) for (int i = 0; i < 5; ++i) dstats[blockIdx.x * 8 + i] = tp[i];
    }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  pdl_trigger();                          // PDL: late on purpose — an early trigger lets grid-stride dependents (LayerNorm, token
                                          // reductions) pile onto the SMs that finish first instead of spreading over all of them
  ptx::cluster_sync();                    // the peer may still be multicasting into / reading from this CTA
  if (warp == 2) ptx::tmem_dealloc_2sm<512>(tmem_base);
}

}  // namespace peft
