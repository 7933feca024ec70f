// C-ABI: LayerNorm, attention, patch-embedding helpers and the skinny adapter-gradient reductions.
#include <stdlib.h>
#include "common.cuh"
#include "ln_sm100.cuh"
#include "attention_sm100.cuh"
#include "attention2_sm100.cuh"
#include "attention2_bwd_sm100.cuh"
#include "attention3_bwd_sm100.cuh"
#include "reduce_sm100.cuh"

using namespace peft;

namespace {
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
inline int ln_grid(int M, bool param_grad = false) {
  const int blocks = (M + 7) / 8;
  // with dgamma / dbeta every CTA ends in a shared-memory reduction and 2*D global atomics: measured on B200 the kernel
  // gets monotonically faster with fewer CTAs (M = 12800: 62.6 us at 8 per SM, 48.1 at 4, 35.9 at 1), so one per SM
  static int pg_cap = 0;                       // tuning knob (tests/bench_ops.py ln_pg): PEFT_LN_PG_CAP CTAs per SM
  if (pg_cap == 0) {
    const char* e = getenv("PEFT_LN_PG_CAP");
    pg_cap = (e && atoi(e) > 0) ? atoi(e) : 1;
  }
  const int cap = sm_count() * (param_grad ? pg_cap : 8);
  return blocks < cap ? (blocks > 0 ? blocks : 1) : cap;
}
}  // namespace

#define LN_DISPATCH_NV(NVVAL, ...)                         \
  switch (NVVAL) {                                         \
    case 1: { constexpr int NV = 1; __VA_ARGS__; break; }  \
    case 2: { constexpr int NV = 2; __VA_ARGS__; break; }  \
    case 3: { constexpr int NV = 3; __VA_ARGS__; break; }  \
    case 4: { constexpr int NV = 4; __VA_ARGS__; break; }  \
    case 6: { constexpr int NV = 6; __VA_ARGS__; break; }  \
    case 8: { constexpr int NV = 8; __VA_ARGS__; break; }  \
    case 10: { constexpr int NV = 10; __VA_ARGS__; break; } \
    default: return fail(PEFT_ERR_UNSUPPORTED, "LayerNorm width %d unsupported (need D/128 in {1,2,3,4,6,8,10})", (NVVAL) * 128); \
  }

extern "C" int peft_layernorm_fwd(const void* x, long long ldx, int x_is_bf16, const float* gamma, const float* beta,
                                  void* y, long long ldy, int y_is_f32, float* mean, float* rstd, int M, int D,
                                  float eps, void* stream) {
  PEFT_REQUIRE(M > 0 && D > 0 && D % 128 == 0, "peft_layernorm_fwd: bad shape M=%d D=%d", M, D);
  PEFT_REQUIRE(aligned16(x) && aligned16(y) && aligned16(gamma) && aligned16(beta), "peft_layernorm_fwd: unaligned pointer");
  PEFT_REQUIRE(ldx % 4 == 0 && ldy % 4 == 0, "peft_layernorm_fwd: pitches must be multiples of 4 elements");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = ln_grid(M);
  LN_DISPATCH_NV(D / 128, {
    if (!x_is_bf16 && !y_is_f32)
      (void)launch_pdl(PDL_LN, ln_fwd_kernel<NV, float, __nv_bfloat16>, dim3(grid), dim3(256), 0, s, (const float*)x, ldx, gamma, beta, (__nv_bfloat16*)y, ldy, mean, rstd, M, eps, nullptr, 0, nullptr, 0);
    else if (!x_is_bf16 && y_is_f32)
      (void)launch_pdl(PDL_LN, ln_fwd_kernel<NV, float, float>, dim3(grid), dim3(256), 0, s, (const float*)x, ldx, gamma, beta, (float*)y, ldy, mean, rstd, M, eps, nullptr, 0, nullptr, 0);
    else if (x_is_bf16 && !y_is_f32)
      (void)launch_pdl(PDL_LN, ln_fwd_kernel<NV, __nv_bfloat16, __nv_bfloat16>, dim3(grid), dim3(256), 0, s, (const __nv_bfloat16*)x, ldx, gamma, beta, (__nv_bfloat16*)y, ldy, mean, rstd, M, eps, nullptr, 0, nullptr, 0);
    else
      (void)launch_pdl(PDL_LN, ln_fwd_kernel<NV, __nv_bfloat16, float>, dim3(grid), dim3(256), 0, s, (const __nv_bfloat16*)x, ldx, gamma, beta, (float*)y, ldy, mean, rstd, M, eps, nullptr, 0, nullptr, 0);
  });
  PEFT_LAUNCH_CHECK();
  return 0;
}

// y(bf16) = LN(x), and add_out(fp32) = x + add_in(fp32) in the same pass.  replaces: the adapter's own LayerNorm plus the
// two residual adds of `x + adapter(mlp(ln_2(x)))` (FS/adapter_model.py:330-334, FS/compacter_model.py:477-479).
extern "C" int peft_layernorm_fwd_add(const float* x, long long ldx, const float* gamma, const float* beta, void* y,
                                      long long ldy, float* mean, float* rstd, const float* add_in, long long ldai,
                                      float* add_out, long long ldao, int M, int D, float eps, void* stream) {
  PEFT_REQUIRE(M > 0 && D > 0 && D % 128 == 0, "peft_layernorm_fwd_add: bad shape M=%d D=%d", M, D);
  PEFT_REQUIRE(aligned16(x) && aligned16(y) && aligned16(gamma) && aligned16(beta) && add_in && add_out && aligned16(add_in) &&
                   aligned16(add_out),
               "peft_layernorm_fwd_add: null or unaligned pointer");
  PEFT_REQUIRE(ldx % 4 == 0 && ldy % 4 == 0 && ldai % 4 == 0 && ldao % 4 == 0, "peft_layernorm_fwd_add: pitches must be multiples of 4");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = ln_grid(M);
  LN_DISPATCH_NV(D / 128, {
    (void)launch_pdl(PDL_LN, ln_fwd_kernel<NV, float, __nv_bfloat16>, dim3(grid), dim3(256), 0, s, x, ldx, gamma, beta, (__nv_bfloat16*)y, ldy, mean, rstd, M, eps,
                                                               add_in, ldai, add_out, ldao);
  });
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_layernorm_bwd(const void* dy, long long lddy, int dy_is_f32, const void* x, long long ldx,
                                  int x_is_bf16, const float* mean, const float* rstd, const float* gamma,
                                  const void* g_in, long long ldgi, void* g_out, long long ldgo, float* dgamma,
                                  float* dbeta, int M, int D, void* stream) {
  PEFT_REQUIRE(M > 0 && D > 0 && D % 128 == 0, "peft_layernorm_bwd: bad shape M=%d D=%d", M, D);
  PEFT_REQUIRE(aligned16(dy) && aligned16(x) && aligned16(g_out) && aligned16(gamma) && (g_in == nullptr || aligned16(g_in)),
               "peft_layernorm_bwd: unaligned pointer");
  PEFT_REQUIRE(lddy % 4 == 0 && ldx % 4 == 0 && ldgo % 4 == 0 && ldgi % 4 == 0, "peft_layernorm_bwd: pitches must be multiples of 4");
  PEFT_REQUIRE((dgamma == nullptr) == (dbeta == nullptr), "peft_layernorm_bwd: dgamma and dbeta go together");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const bool pg = dgamma != nullptr;
  const int grid = ln_grid(M, pg);
  using bf = __nv_bfloat16;
#define LN_BWD_LAUNCH(DYT, XT, PG) \
  (void)launch_pdl(PDL_LN, ln_bwd_kernel<NV, DYT, XT, bf, PG>, dim3(grid), dim3(256), 0, s, (const DYT*)dy, lddy, (const XT*)x, ldx, mean, rstd, gamma, (const bf*)g_in, ldgi, (bf*)g_out, ldgo, dgamma, dbeta, M, nullptr, 0)
  LN_DISPATCH_NV(D / 128, {
    if (dy_is_f32) {
      if (x_is_bf16) { if (pg) LN_BWD_LAUNCH(float, bf, true); else LN_BWD_LAUNCH(float, bf, false); }
      else { if (pg) LN_BWD_LAUNCH(float, float, true); else LN_BWD_LAUNCH(float, float, false); }
    } else {
      if (x_is_bf16) { if (pg) LN_BWD_LAUNCH(bf, bf, true); else LN_BWD_LAUNCH(bf, bf, false); }
      else { if (pg) LN_BWD_LAUNCH(bf, float, true); else LN_BWD_LAUNCH(bf, float, false); }
    }
  });
#undef LN_BWD_LAUNCH
  PEFT_LAUNCH_CHECK();
  return 0;
}

// fp32 residual-gradient stream variant: g_in / g_out are fp32 (may alias), g_copy (optional) receives the bf16 copy the
// next dX GEMM consumes as its A operand.  Keeps the 2 x depth residual-gradient adds out of bf16 (precision).
extern "C" int peft_layernorm_bwd_f32g(const void* dy, long long lddy, int dy_is_f32, const void* x, long long ldx,
                                       int x_is_bf16, const float* mean, const float* rstd, const float* gamma,
                                       const float* g_in, long long ldgi, float* g_out, long long ldgo, void* g_copy,
                                       long long ldgc, float* dgamma, float* dbeta, int M, int D, void* stream) {
  PEFT_REQUIRE(M > 0 && D > 0 && D % 128 == 0, "peft_layernorm_bwd_f32g: bad shape M=%d D=%d", M, D);
  PEFT_REQUIRE(aligned16(dy) && aligned16(x) && aligned16(g_out) && aligned16(gamma) && (g_in == nullptr || aligned16(g_in)) &&
                   (g_copy == nullptr || aligned16(g_copy)),
               "peft_layernorm_bwd_f32g: unaligned pointer");
  PEFT_REQUIRE(lddy % 4 == 0 && ldx % 4 == 0 && ldgo % 4 == 0 && ldgi % 4 == 0 && ldgc % 4 == 0,
               "peft_layernorm_bwd_f32g: pitches must be multiples of 4");
  PEFT_REQUIRE((dgamma == nullptr) == (dbeta == nullptr), "peft_layernorm_bwd_f32g: dgamma and dbeta go together");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const bool pg = dgamma != nullptr;
  const int grid = ln_grid(M, pg);
  using bf = __nv_bfloat16;
#define LN_BWD_LAUNCH32(DYT, XT, PG) \
  (void)launch_pdl(PDL_LN, ln_bwd_kernel<NV, DYT, XT, float, PG>, dim3(grid), dim3(256), 0, s, (const DYT*)dy, lddy, (const XT*)x, ldx, mean, rstd, gamma, g_in, ldgi, g_out, ldgo, dgamma, dbeta, M, (bf*)g_copy, ldgc)
  LN_DISPATCH_NV(D / 128, {
    if (dy_is_f32) {
      if (x_is_bf16) { if (pg) LN_BWD_LAUNCH32(float, bf, true); else LN_BWD_LAUNCH32(float, bf, false); }
      else { if (pg) LN_BWD_LAUNCH32(float, float, true); else LN_BWD_LAUNCH32(float, float, false); }
    } else {
      if (x_is_bf16) { if (pg) LN_BWD_LAUNCH32(bf, bf, true); else LN_BWD_LAUNCH32(bf, bf, false); }
      else { if (pg) LN_BWD_LAUNCH32(bf, float, true); else LN_BWD_LAUNCH32(bf, float, false); }
    }
  });
#undef LN_BWD_LAUNCH32
  PEFT_LAUNCH_CHECK();
  return 0;
}

static int attn_lp(int L) { return ((L + 63) / 64) * 64; }
static int attn_prefetch() {
  // measured r02 (bench_ops.py attn, B = 256, H = 12, L = 197): no gain (fwd 84.8 -> 87.6 us, bwd 330 -> 337 us), so it is OFF by
  // default and kept as a switch: the kernels are not bound by the latency of their operand loads
  static const int v = [] { const char* e = getenv("PEFT_ATTN_PREFETCH"); return e ? atoi(e) : 0; }();
  return v;
}

static int attn_check(const AttnParams& p) {
  PEFT_REQUIRE(p.B > 0 && p.H > 0 && p.L > 0, "attention: empty problem");
  PEFT_REQUIRE(p.D == p.H * kHd, "attention: head_dim must be 64 (D=%d, H=%d)", p.D, p.H);
  PEFT_REQUIRE(p.L <= 512, "attention: L=%d exceeds the single-head-resident limit (512)", p.L);
  PEFT_REQUIRE(p.ldqkv % 8 == 0 && aligned16(p.qkv), "attention: qkv must be 16-byte aligned with pitch % 8 == 0");
  return 0;
}

static std::atomic<int>& attn_algo() {
  // 0 = auto (tcgen05 kernels, pipelined backward), 1 = mma.sync v1 kernels, 2 = tcgen05 kernels with the round-1 backward
  // (attention2_bwd_sm100.cuh: tensor core and exp pipes take turns), 3 = same as auto
  static std::atomic<int> a{[] { const char* e = getenv("PEFT_ATTN_ALGO"); const int v = e ? atoi(e) : 0; return (v >= 0 && v <= 3) ? v : 0; }()};
  return a;
}
extern "C" int peft_set_attention_algo(int algo) {
  PEFT_REQUIRE(algo >= 0 && algo <= 3, "peft_set_attention_algo: algo must be 0 (auto), 1 (mma.sync), 2 (tcgen05, unpipelined backward) or 3 (tcgen05, pipelined backward)");
  attn_algo().store(algo);
  return 0;
}

// diagnostics: per-phase cycle counters of the LAST tcgen05 attention-forward launch (env PEFT_ATTN_STATS=1)
static unsigned long long* attn_stats_buffer() {
  static unsigned long long* buf = [] {
    const char* e = getenv("PEFT_ATTN_STATS");
    unsigned long long* p = nullptr;
    if (e && atoi(e) > 0 && cudaMalloc(&p, 512 * 16 * sizeof(unsigned long long)) == cudaSuccess) cudaMemset(p, 0, 512 * 16 * sizeof(unsigned long long));
    else p = nullptr;
    return p;
  }();
  return buf;
}
extern "C" int peft_debug_attn_stats(unsigned long long* host_out, int n_ctas) {
#if !PEFT_GEMM_DIAG
  return fail(PEFT_ERR_INVALID, "peft_debug_attn_stats: the library was built without the in-kernel diagnostics (make DIAG=1)");
#endif
  unsigned long long* buf = attn_stats_buffer();
  PEFT_REQUIRE(buf != nullptr, "peft_debug_attn_stats: set PEFT_ATTN_STATS=1 before the first attention call");
  PEFT_REQUIRE(host_out && n_ctas > 0 && n_ctas <= 512, "peft_debug_attn_stats: bad arguments");
  PEFT_CUDA_CHECK(cudaDeviceSynchronize());
  PEFT_CUDA_CHECK(cudaMemcpy(host_out, buf, (size_t)n_ctas * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return 0;
}

template <int NK>
static int launch_attn2_fwd(const void* qkv, long long ldqkv, void* o, long long ldo, float* lse2, int B, int H, int L,
                            float scale, cudaStream_t s) {
  using C = Attn2Cfg<NK>;
  CUtensorMap tq, tkv, to;
  int rc = make_tmap_bf16_3d(&tq, qkv, B, L, 3ull * H * 64, ldqkv, 128);
  if (rc) return rc;
  rc = make_tmap_bf16_3d(&tkv, qkv, B, L, 3ull * H * 64, ldqkv, C::kKVBox);
  if (rc) return rc;
  rc = make_tmap_bf16_3d(&to, o, B, L, 1ull * H * 64, ldo, 128);
  if (rc) return rc;
  auto kern = attn2_fwd_kernel<NK>;
  static DeviceOnce configured;
  if (configured.needs()) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    configured.done();
  }
  Attn2Params p{};
  p.lse2 = lse2; p.B = B; p.H = H; p.L = L; p.scale = scale;
  p.stats = attn_stats_buffer();
  const int grid = B * H < sm_count() ? B * H : sm_count();
  (void)launch_pdl(PDL_ATTN, kern, dim3(grid), dim3(kA2Threads), C::kSmemBytes, s, tq, tkv, to, p);
  PEFT_LAUNCH_CHECK_KIND(KIND_ATTN_TC_FWD);
  return 0;
}

extern "C" int peft_attention_fwd(const void* qkv, long long ldqkv, void* o, long long ldo, float* lse2, int B, int H,
                                  int L, float scale, void* stream) {
  AttnParams p{};
  p.qkv = (const __nv_bfloat16*)qkv; p.ldqkv = ldqkv; p.o = (__nv_bfloat16*)o; p.ldo = ldo; p.lse2 = lse2;
  p.B = B; p.H = H; p.L = L; p.D = H * kHd; p.scale = scale;
  int rc = attn_check(p);
  if (rc) return rc;
  PEFT_REQUIRE(o && lse2 && ldo % 8 == 0 && aligned16(o), "peft_attention_fwd: bad output");
  if (attn_algo().load() != 1 && L <= 272) {
    cudaStream_t s2 = reinterpret_cast<cudaStream_t>(stream);
    if (L <= 64) return launch_attn2_fwd<64>(qkv, ldqkv, o, ldo, lse2, B, H, L, scale, s2);
    if (L <= 208) return launch_attn2_fwd<208>(qkv, ldqkv, o, ldo, lse2, B, H, L, scale, s2);
    return launch_attn2_fwd<272>(qkv, ldqkv, o, ldo, lse2, B, H, L, scale, s2);
  }
  const int Lp = attn_lp(L);
  const int smem = (64 + 2 * Lp) * 128;
  static DeviceMax configured;
  if (configured.get() < smem) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured.set(smem);
  }
  dim3 grid((L + 63) / 64, B * H);
  attn_fwd_kernel<<<grid, kAttnThreads, smem, reinterpret_cast<cudaStream_t>(stream)>>>(p, Lp);
  PEFT_LAUNCH_CHECK_KIND(KIND_ATTN_LEGACY);
  return 0;
}

template <int NKC, int NCH>
static int launch_attn2_bwd(const void* qkv, long long ldqkv, const void* o, long long ldo, const void* d_o, long long lddo,
                            const float* lse2, float* delta, void* dqkv, long long lddqkv, int B, int H, int L, float scale,
                            cudaStream_t s) {
  using C = Attn2BwdCfg<NKC, NCH>;
  const unsigned long long D = 64ull * H;
  CUtensorMap t128, tchunk, tdo128, tdochunk, to128, tout;
  int rc;
  if ((rc = make_tmap_bf16_3d(&t128, qkv, B, L, 3 * D, ldqkv, 128))) return rc;
  if ((rc = make_tmap_bf16_3d(&tchunk, qkv, B, L, 3 * D, ldqkv, NKC))) return rc;
  if ((rc = make_tmap_bf16_3d(&tdo128, d_o, B, L, D, lddo, 128))) return rc;
  if ((rc = make_tmap_bf16_3d(&tdochunk, d_o, B, L, D, lddo, NKC))) return rc;
  if ((rc = make_tmap_bf16_3d(&to128, o, B, L, D, ldo, 128))) return rc;
  if ((rc = make_tmap_bf16_3d(&tout, dqkv, B, L, 3 * D, lddqkv, 128))) return rc;
  auto kdq = attn2_bwd_dq_kernel<NKC, NCH>;
  auto kdkv = attn2_bwd_dkv_kernel<NKC, NCH>;
  static DeviceOnce configured;
  if (configured.needs()) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(kdq, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kDqSmemBytes));
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(kdkv, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kDkvSmemBytes));
    configured.done();
  }
  Attn2Params p{};
  p.lse2 = const_cast<float*>(lse2); p.delta_in = delta; p.delta_out = delta; p.B = B; p.H = H; p.L = L; p.scale = scale;
  const int grid = B * H < sm_count() ? B * H : sm_count();
  (void)launch_pdl(PDL_ATTN, kdq, dim3(grid), dim3(kA2Threads), C::kDqSmemBytes, s, t128, tchunk, tdo128, to128, tout, p);
  PEFT_LAUNCH_CHECK_KIND(KIND_ATTN_TC_BWD);
  (void)launch_pdl(PDL_ATTN, kdkv, dim3(grid), dim3(kA2Threads), C::kDkvSmemBytes, s, t128, tchunk, tdochunk, tout, p);
  PEFT_LAUNCH_CHECK_KIND(KIND_ATTN_TC_BWD);
  return 0;
}

// software-pipelined backward (attention3_bwd_sm100.cuh)
template <int NKT>
static int launch_attn3_bwd(const void* qkv, long long ldqkv, const void* o, long long ldo, const void* d_o, long long lddo,
                            const float* lse2, float* delta, void* dqkv, long long lddqkv, int B, int H, int L, float scale,
                            cudaStream_t s) {
  using C = Attn3Cfg<NKT>;
  const unsigned long long D = 64ull * H;
  CUtensorMap t128, thead, tdo128, tdohead, to128;
  int rc;
  if ((rc = make_tmap_bf16_3d(&t128, qkv, B, L, 3 * D, ldqkv, 128))) return rc;
  if ((rc = make_tmap_bf16_3d(&thead, qkv, B, L, 3 * D, ldqkv, C::kBox))) return rc;
  if ((rc = make_tmap_bf16_3d(&tdo128, d_o, B, L, D, lddo, 128))) return rc;
  if ((rc = make_tmap_bf16_3d(&tdohead, d_o, B, L, D, lddo, C::kBox))) return rc;
  if ((rc = make_tmap_bf16_3d(&to128, o, B, L, D, ldo, 128))) return rc;
  auto kdq = attn3_bwd_dq_kernel<NKT>;
  auto kdkv = attn3_bwd_dkv_kernel<NKT>;
  static DeviceOnce configured;
  if (configured.needs()) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(kdq, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kDqSmemBytes));
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(kdkv, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kDkvSmemBytes));
    configured.done();
  }
  Attn3Params p{};
  p.lse2 = lse2; p.delta_in = delta; p.delta_out = delta; p.dqkv = (__nv_bfloat16*)dqkv; p.lddqkv = lddqkv;
  p.B = B; p.H = H; p.L = L; p.scale = scale; p.prefetch = attn_prefetch();
  p.stats = attn_stats_buffer();
  static const int a3_flags = [] { const char* e = getenv("PEFT_ATTN3_FLAGS"); return e ? atoi(e) : 0; }();
  p.flags = a3_flags;
  const int grid = B * H < sm_count() ? B * H : sm_count();
  (void)launch_pdl(PDL_ATTN, kdq, dim3(grid), dim3(kA2Threads), C::kDqSmemBytes, s, t128, thead, tdo128, to128, p);
  PEFT_LAUNCH_CHECK_KIND(KIND_ATTN_TC_BWD);
  (void)launch_pdl(PDL_ATTN, kdkv, dim3(grid), dim3(kA2Threads), C::kDkvSmemBytes, s, t128, thead, tdohead, p);
  PEFT_LAUNCH_CHECK_KIND(KIND_ATTN_TC_BWD);
  return 0;
}

extern "C" int peft_attention_bwd(const void* qkv, long long ldqkv, const void* o, long long ldo, const void* d_o,
                                  long long lddo, const float* lse2, float* delta, void* dqkv, long long lddqkv, int B,
                                  int H, int L, float scale, void* stream) {
  AttnParams p{};
  p.qkv = (const __nv_bfloat16*)qkv; p.ldqkv = ldqkv; p.o = (__nv_bfloat16*)const_cast<void*>(o); p.ldo = ldo;
  p.d_o = (const __nv_bfloat16*)d_o; p.lddo = lddo; p.dqkv = (__nv_bfloat16*)dqkv; p.lddqkv = lddqkv;
  p.lse2 = const_cast<float*>(lse2); p.delta = delta;
  p.B = B; p.H = H; p.L = L; p.D = H * kHd; p.scale = scale;
  int rc = attn_check(p);
  if (rc) return rc;
  PEFT_REQUIRE(o && d_o && lse2 && delta && dqkv, "peft_attention_bwd: null operand");
  PEFT_REQUIRE(ldo % 8 == 0 && lddo % 8 == 0 && lddqkv % 8 == 0 && aligned16(o) && aligned16(d_o) && aligned16(dqkv),
               "peft_attention_bwd: operands must be 16-byte aligned with pitch % 8 == 0");
  // pipelined backward: measured faster for one-tile heads (L <= 64: 98 vs 106 us at B = 256, H = 12, L = 50) and slower for the
  // longer ones (L = 197: 432 vs 331 us) — there both variants are bound by the delivery rate of the head-strided operand rows,
  // not by the tensor / exp overlap it adds — so automatic mode takes it for L <= 64 only; algo 3 forces it
  const int algo = attn_algo().load();
  if ((algo == 3 || (algo == 0 && L <= 64)) && L <= 272) {
    cudaStream_t s2 = reinterpret_cast<cudaStream_t>(stream);
    if (L <= 64) return launch_attn3_bwd<64>(qkv, ldqkv, o, ldo, d_o, lddo, lse2, delta, dqkv, lddqkv, B, H, L, scale, s2);
    if (L <= 208) return launch_attn3_bwd<208>(qkv, ldqkv, o, ldo, d_o, lddo, lse2, delta, dqkv, lddqkv, B, H, L, scale, s2);
    return launch_attn3_bwd<272>(qkv, ldqkv, o, ldo, d_o, lddo, lse2, delta, dqkv, lddqkv, B, H, L, scale, s2);
  }
  if (attn_algo().load() != 1 && L <= 288) {
    cudaStream_t s2 = reinterpret_cast<cudaStream_t>(stream);
    if (L <= 64) return launch_attn2_bwd<64, 1>(qkv, ldqkv, o, ldo, d_o, lddo, lse2, delta, dqkv, lddqkv, B, H, L, scale, s2);
    if (L <= 208) return launch_attn2_bwd<208, 1>(qkv, ldqkv, o, ldo, d_o, lddo, lse2, delta, dqkv, lddqkv, B, H, L, scale, s2);
    return launch_attn2_bwd<144, 2>(qkv, ldqkv, o, ldo, d_o, lddo, lse2, delta, dqkv, lddqkv, B, H, L, scale, s2);
  }
  const int Lp = attn_lp(L);
  const int smem_dq = (128 + 2 * Lp) * 128;
  const int smem_dkv = (128 + 2 * Lp) * 128 + 2 * Lp * 4;
  static DeviceMax conf_dq, conf_dkv;
  if (conf_dq.get() < smem_dq) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dq));
    conf_dq.set(smem_dq);
  }
  if (conf_dkv.get() < smem_dkv) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dkv));
    conf_dkv.set(smem_dkv);
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid((L + 63) / 64, B * H);
  attn_bwd_dq_kernel<<<grid, kAttnThreads, smem_dq, s>>>(p, Lp);
  PEFT_LAUNCH_CHECK_KIND(KIND_ATTN_LEGACY);
  attn_bwd_dkv_kernel<<<grid, kAttnThreads, smem_dkv, s>>>(p, Lp);
  PEFT_LAUNCH_CHECK_KIND(KIND_ATTN_LEGACY);
  return 0;
}

extern "C" int peft_im2col_patches(const float* images, void* patches, long long ldp, int B, int img, int patch, void* stream) {
  PEFT_REQUIRE(B > 0 && img > 0 && patch > 0 && img % patch == 0 && patch % 2 == 0, "peft_im2col_patches: bad shape (img=%d patch=%d)", img, patch);
  PEFT_REQUIRE(aligned16(images) && aligned16(patches) && ldp >= 3ll * patch * patch && ldp % 8 == 0,
               "peft_im2col_patches: unaligned pointer or pitch (need ldp %% 8 == 0, ldp >= 3*patch*patch)");
  const int vec = (patch % 4 == 0) ? 4 : 2;
  const long long total = (long long)B * (img / patch) * (img / patch) * 3 * patch * patch / vec;
  long long blocks = (total + 255) / 256;
  if (blocks > sm_count() * 16) blocks = sm_count() * 16;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (vec == 4) im2col_patch_kernel<4><<<(int)blocks, 256, 0, s>>>(images, (__nv_bfloat16*)patches, ldp, B, img, patch);
  else im2col_patch_kernel<2><<<(int)blocks, 256, 0, s>>>(images, (__nv_bfloat16*)patches, ldp, B, img, patch);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_assemble_tokens(const void* patch_embed, const float* cls, const float* pos, float* x, int B, int L,
                                    int D, void* stream) {
  PEFT_REQUIRE(B > 0 && L > 1 && D % 4 == 0, "peft_assemble_tokens: bad shape");
  PEFT_REQUIRE(aligned16(patch_embed) && aligned16(cls) && aligned16(pos) && aligned16(x), "peft_assemble_tokens: unaligned pointer");
  const long long total = (long long)B * L * D / 4;
  long long blocks = (total + 255) / 256;
  if (blocks > sm_count() * 16) blocks = sm_count() * 16;
  assemble_tokens_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>((const __nv_bfloat16*)patch_embed, cls, pos, x, B, L, D);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_tn_reduce(const void* x, long long ldx, int I, const void* y, long long ldy, int J, float* out,
                              long long ldout, int M, void* stream) {
  PEFT_REQUIRE(M > 0 && (I == 16 || I == 32 || I == 64) && J > 0 && J % 64 == 0, "peft_tn_reduce: need I in {16,32,64}, J %% 64 == 0 (I=%d J=%d)", I, J);
  PEFT_REQUIRE(aligned16(x) && aligned16(y) && ldx % 8 == 0 && ldy % 8 == 0, "peft_tn_reduce: operands must be 16-byte aligned with pitch %% 8 == 0");
  const int jt = J / 64;
  // every CTA ends in I x 64 global atomics: few, long CTAs (>= 4 stages of 64 rows each); long reductions take four CTAs per
  // SM so that enough cp.async stages are in flight to cover the DRAM latency
  // (measured r02 at M = 50 432, J = 2304 / 768: x3 65.9 / 27.8 us, x4 68.4 / 25.1, x6 61.7 / 29.0, x8 65.8 / 30.3; two-stage ring
  // with x2, the round-1 form: 103.5 / 41.9 us)
  static const int red_mult = [] { const char* e = getenv("PEFT_RED_CTAS"); return e ? atoi(e) : 6; }();
  int target_ctas = sm_count() * (M >= 16384 ? red_mult : 2);
  int my = (target_ctas + jt - 1) / jt;
  int rows = (M + my - 1) / my;
  rows = ((rows + 63) / 64) * 64;
  if (rows < 256) rows = 256;
  my = (M + rows - 1) / rows;
  dim3 grid(jt, my);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const __nv_bfloat16* X = (const __nv_bfloat16*)x; const __nv_bfloat16* Y = (const __nv_bfloat16*)y;
  if (I == 16) (void)launch_pdl(PDL_REDUCE, tn_reduce_kernel<1>, dim3(grid), dim3(kRedThreads), 0, s, X, ldx, Y, ldy, out, ldout, M, rows);
  else if (I == 32) (void)launch_pdl(PDL_REDUCE, tn_reduce_kernel<2>, dim3(grid), dim3(kRedThreads), 0, s, X, ldx, Y, ldy, out, ldout, M, rows);
  else (void)launch_pdl(PDL_REDUCE, tn_reduce_kernel<4>, dim3(grid), dim3(kRedThreads), 0, s, X, ldx, Y, ldy, out, ldout, M, rows);
  PEFT_LAUNCH_CHECK();
  return 0;
}

// diagnostics: cycles of a chain of `reps` tcgen05.mma (M = 128, N = n, K = 16) rotating over `nacc` accumulators; out[0] = issue
// cycles, out[1] = cycles until the last one has completed (host memory, 2 values)
extern "C" int peft_debug_mma_probe(int n, int nacc, int ts, int reps, unsigned long long* host_out) {
  PEFT_REQUIRE(n >= 16 && n <= 256 && n % 16 == 0 && nacc >= 1 && nacc * n <= 256 && reps > 0 && host_out, "peft_debug_mma_probe: bad arguments");
  unsigned long long* d = nullptr;
  PEFT_CUDA_CHECK(cudaMalloc(&d, 16));
  const int smem = 2 * 32768 + 1024;
  PEFT_REQUIRE(nacc == 1 || nacc == 2 || nacc == 4, "peft_debug_mma_probe: nacc must be 1, 2 or 4");
  auto kern = nacc == 1 ? mma_probe_kernel<1> : (nacc == 2 ? mma_probe_kernel<2> : mma_probe_kernel<4>);
  PEFT_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<1, 128, smem>>>(n, ts, reps, d);
  PEFT_CUDA_CHECK(cudaDeviceSynchronize());
  PEFT_CUDA_CHECK(cudaMemcpy(host_out, d, 16, cudaMemcpyDeviceToHost));
  cudaFree(d);
  return 0;
}
