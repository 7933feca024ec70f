// C-ABI: peft_gemm_bf16 — host launcher of the tcgen05/TMA GEMM (gemm_sm100.cuh).
#include <stdlib.h>
#include "common.cuh"
#include "gemm_sm100.cuh"
#include "gemm2_sm100.cuh"

namespace peft {

template <int BN, int EPI>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kern = gemm_bf16_tcgen05_kernel<BN, EPI>;
  static DeviceOnce configured;  // per (BN,EPI) instantiation
  if (configured.needs()) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured.done();
  }
  const int num_tiles = ((p.M + kBM - 1) / kBM) * ((p.N + BN - 1) / BN);
  const int grid = num_tiles < sm_count() ? num_tiles : sm_count();
  (void)launch_pdl(PDL_GEMM1, kern, dim3(grid), dim3(kGemmThreads), Cfg::kSmemBytes, stream, ta, tb, p);
  PEFT_LAUNCH_CHECK_KIND(KIND_GEMM_1CTA);
  return 0;
}

template <int BN>
static int dispatch_epi(int epi, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t s) {
  switch (epi) {
    case EPI_STORE: return launch_gemm<BN, EPI_STORE>(ta, tb, p, s);
    case EPI_GELU: return launch_gemm<BN, EPI_GELU>(ta, tb, p, s);
    case EPI_QUICKGELU: return launch_gemm<BN, EPI_QUICKGELU>(ta, tb, p, s);
    case EPI_RESID_F32: return launch_gemm<BN, EPI_RESID_F32>(ta, tb, p, s);
    case EPI_DGELU: return launch_gemm<BN, EPI_DGELU>(ta, tb, p, s);
    case EPI_DQUICKGELU: return launch_gemm<BN, EPI_DQUICKGELU>(ta, tb, p, s);
    case EPI_STORE_F32: return launch_gemm<BN, EPI_STORE_F32>(ta, tb, p, s);
    case EPI_RELU: return launch_gemm<BN, EPI_RELU>(ta, tb, p, s);
    case EPI_DRELU: return launch_gemm<BN, EPI_DRELU>(ta, tb, p, s);
    case EPI_GELU_TANH: return launch_gemm<BN, EPI_GELU_TANH>(ta, tb, p, s);
    case EPI_DGELU_TANH: return launch_gemm<BN, EPI_DGELU_TANH>(ta, tb, p, s);
    case EPI_RESID_BF16: return launch_gemm<BN, EPI_RESID_BF16>(ta, tb, p, s);
    default: return fail(PEFT_ERR_INVALID, "unknown epilogue %d", epi);
  }
}

// diagnostics: per-CTA wait-cycle counters of the LAST 2-CTA GEMM launch (env PEFT_GEMM_STATS=1 allocates the buffer)
static unsigned long long* gemm_stats_buffer() {
  static unsigned long long* buf = [] {
    const char* e = getenv("PEFT_GEMM_STATS");
    unsigned long long* p = nullptr;
    if (e && atoi(e) > 0 && cudaMalloc(&p, 256 * 8 * sizeof(unsigned long long)) == cudaSuccess) cudaMemset(p, 0, 256 * 8 * sizeof(unsigned long long));
    else p = nullptr;
    return p;
  }();
  return buf;
}

// ---------------------------------------------------------------- stream-K (see GemmParams::sk_dp_tiles, TileSched)
static std::atomic<int>& streamk_mode() {
  static std::atomic<int> m{[] { const char* e = getenv("PEFT_GEMM_STREAMK"); return e ? atoi(e) : 1; }()};
  return m;       // 0 = off, 1 = automatic, 2 = whenever the shape allows it (tests)
}
// One workspace per device, allocated at the first 2-CTA launch (i.e. during warm-up, never inside a graph capture): a
// [256 x 256] fp32 slot per cluster + per-warp flags.  Stream-K launches of one device must not overlap in time (they are
// issued on one stream by the engine); concurrent multi-stream callers switch it off with peft_set_gemm_streamk(0).
struct SkWorkspace { float4* ws = nullptr; int* flags = nullptr; bool tried = false; };
static SkWorkspace* sk_workspace() {
  static std::mutex mu;
  static SkWorkspace per_dev[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> g(mu);
  SkWorkspace& w = per_dev[dev];
  if (!w.tried) {
    w.tried = true;
    const size_t clusters = 128, slot = (size_t)2 * kEpiWarps * 1024 * sizeof(float4);
    void* a = nullptr;
    void* f = nullptr;
    if (cudaMalloc(&a, clusters * slot) == cudaSuccess && cudaMalloc(&f, clusters * 2 * kEpiWarps * sizeof(int)) == cudaSuccess &&
        cudaMemset(f, 0, clusters * 2 * kEpiWarps * sizeof(int)) == cudaSuccess) {
      w.ws = reinterpret_cast<float4*>(a);
      w.flags = reinterpret_cast<int*>(f);
    } else {
      (void)cudaGetLastError();
    }
  }
  return w.ws ? &w : nullptr;
}
// Decide whether (and from which tile on) the product runs stream-K: only long contractions (>= 16 k-blocks: the fix-up
// moves a 256 KB partial tile through L2, which a short K cannot amortise) whose tile count leaves the last wave mostly
// empty.  All but the last full wave stay data-parallel; the last full wave + the remainder are dealt out by k-blocks.
static void decide_streamk(GemmParams& p, int M, int N, int K) {
  p.sk_dp_tiles = -1;
  const int mode = streamk_mode().load(std::memory_order_relaxed);
  if (mode == 0 || p.ksplit > 1) return;
  SkWorkspace* w = sk_workspace();
  if (w == nullptr) return;
  const long long T = (long long)((M + 255) / 256) * ((N + 255) / 256);
  const int C = sm_count() / 2;
  const int KB = (K + kBK - 1) / kBK;
  if (C > 128 || T > (1 << 24)) return;
  const long long waves = (T + C - 1) / C;
  const double eff = (double)T / (double)(waves * C);
  // the 16 k-block gate re-measured at per-GPU batch 32 (C3 step, graph replay): gate 16 -> 4.91 ms, 12 -> 5.00 ms, 8 -> 5.08 ms
  static const int min_kb = [] { const char* e = getenv("PEFT_GEMM_SK_MINKB"); return e ? atoi(e) : 16; }();
  if (mode == 1 && (KB < min_kb || eff >= 0.92)) return;
  if (mode == 2 && KB < 2) return;
  const long long dp = T >= C ? (T / C - 1) * C : 0;
  if ((T - dp) * KB < 4LL * C) return;          // every cluster must get a non-empty k-block range
  p.sk_dp_tiles = (int)dp;
  p.sk_ws = w->ws;
  p.sk_flags = w->flags;
}

// ---------------------------------------------------------------- 2-CTA kernel (gemm2_sm100.cuh)
static std::atomic<int>& gemm_algo() {
  static std::atomic<int> a{0};   // 0 = auto, 1 = force 1-CTA kernel, 2 = force 2-CTA kernel
  return a;
}

template <int EPI, bool TN = false>
static int launch_gemm2(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& to2,
                        const CUtensorMap& tx, const GemmParams& p, cudaStream_t stream) {
  auto kern = gemm2_bf16_tcgen05_kernel<EPI, TN>;
  static DeviceOnce configured;
  if (configured.needs()) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kG2SmemBytes));
    configured.done();
  }
  const int num_tiles = ((p.M + 2 * kBM - 1) / (2 * kBM)) * ((p.N + kG2BN - 1) / kG2BN) * (p.ksplit > 1 ? p.ksplit : 1);
  const int max_clusters = sm_count() / 2;
  const int clusters = (num_tiles < max_clusters && p.sk_dp_tiles < 0) ? num_tiles : max_clusters;   // stream-K uses every pair
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters, 1, 1);
  cfg.blockDim = dim3(kG2Threads, 1, 1);
  cfg.dynamicSmemBytes = kG2SmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;      // PDL, see common.cuh
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (pdl_mode().load(std::memory_order_relaxed) & PDL_GEMM2) ? 2 : 1;
  PEFT_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, to, to2, tx, p));
  PEFT_LAUNCH_CHECK_KIND(TN ? KIND_GEMM_TN : KIND_GEMM_2CTA);
  return 0;
}

static int dispatch_gemm2(int epi, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to,
                          const CUtensorMap& to2, const CUtensorMap& tx, const GemmParams& p, cudaStream_t s) {
  switch (epi) {
    case EPI_STORE: return launch_gemm2<EPI_STORE>(ta, tb, to, to2, tx, p, s);
    case EPI_GELU: return launch_gemm2<EPI_GELU>(ta, tb, to, to2, tx, p, s);
    case EPI_QUICKGELU: return launch_gemm2<EPI_QUICKGELU>(ta, tb, to, to2, tx, p, s);
    case EPI_RESID_F32: return launch_gemm2<EPI_RESID_F32>(ta, tb, to, to2, tx, p, s);
    case EPI_DGELU: return launch_gemm2<EPI_DGELU>(ta, tb, to, to2, tx, p, s);
    case EPI_DQUICKGELU: return launch_gemm2<EPI_DQUICKGELU>(ta, tb, to, to2, tx, p, s);
    case EPI_STORE_F32: return launch_gemm2<EPI_STORE_F32>(ta, tb, to, to2, tx, p, s);
    case EPI_RELU: return launch_gemm2<EPI_RELU>(ta, tb, to, to2, tx, p, s);
    case EPI_DRELU: return launch_gemm2<EPI_DRELU>(ta, tb, to, to2, tx, p, s);
    case EPI_GELU_TANH: return launch_gemm2<EPI_GELU_TANH>(ta, tb, to, to2, tx, p, s);
    case EPI_DGELU_TANH: return launch_gemm2<EPI_DGELU_TANH>(ta, tb, to, to2, tx, p, s);
    case EPI_RESID_BF16: return launch_gemm2<EPI_RESID_BF16>(ta, tb, to, to2, tx, p, s);
    default: return fail(PEFT_ERR_INVALID, "unknown epilogue %d", epi);
  }
}

// The 2-CTA kernel needs enough 256 x 256 tiles to fill the 74 SM pairs; small / narrow problems stay on the 1-CTA kernel.
static bool use_gemm2(int M, int N) {
  const int a = gemm_algo().load(std::memory_order_relaxed);
  if (a == 1) return false;
  if (a == 2) return true;
  if (N < 256) return false;
  const long long tiles = (long long)((M + 255) / 256) * ((N + 255) / 256);
  return tiles >= sm_count() / 2;
}

// Split-K for weight-gradient-shaped products (fp32 output with few 256 x 256 tiles and a long contraction): the smallest
// number of K ranges (<= 8, each >= 16 k-blocks) that fills >= 90 % of the waves over the 74 SM pairs; 1 = do not split.
static int pick_ksplit(int M, int N, int K, int epilogue, const float* bias, long long ldo) {
  if (epilogue != EPI_STORE_F32 || bias != nullptr || gemm_algo().load(std::memory_order_relaxed) == 1) return 1;
  if (N < 256 || ldo != N) return 1;                      // contiguous output: one memset zeroes it
  const long long tiles = (long long)((M + 255) / 256) * ((N + 255) / 256);
  const int pairs = sm_count() / 2;
  const int num_kb = (K + kBK - 1) / kBK;
  if (tiles >= pairs || num_kb < 32) return 1;
  int best = 1;
  double best_eff = 0.0;
  for (int sp = 2; sp <= 8; ++sp) {
    const int per = (num_kb + sp - 1) / sp;
    if (per < 16 || (long long)per * (sp - 1) >= num_kb) continue;      // no empty range
    const long long t = tiles * sp;
    const double eff = (double)t / (double)(((t + pairs - 1) / pairs) * pairs);
    if (eff > best_eff + 0.02) { best_eff = eff; best = sp; }
    if (best_eff >= 0.9) break;
  }
  return best;
}

// Tile-width heuristic: widest tile that still gives >= ~2 waves over the SMs; narrow outputs get BN=32.
static int pick_bn(int M, int N) {
  if (N <= 64) return 32;
  const long long mt = (M + kBM - 1) / kBM;
  const long long t256 = mt * ((N + 255) / 256);
  if (N % 256 == 0 && t256 >= 2LL * sm_count()) return 256;
  return 128;
}

}  // namespace peft

using namespace peft;

extern "C" int peft_gemm_bf16(const void* a, long long lda, const void* b, long long ldb, int M, int N, int K,
                              const float* bias, void* out, long long ldo, void* out2, long long ldo2,
                              const void* aux, long long ldaux, int epilogue, void* stream) {
  return peft_gemm_bf16_rs(a, lda, b, ldb, M, N, K, bias, out, ldo, out2, ldo2, aux, ldaux, epilogue, nullptr, stream);
}

extern "C" int peft_gemm_bf16_rs(const void* a, long long lda, const void* b, long long ldb, int M, int N, int K,
                                 const float* bias, void* out, long long ldo, void* out2, long long ldo2,
                                 const void* aux, long long ldaux, int epilogue, const float* rowscale, void* stream) {
  PEFT_REQUIRE(M > 0 && N > 0 && K > 0, "peft_gemm_bf16: empty problem M=%d N=%d K=%d", M, N, K);
  PEFT_REQUIRE(a && b && out, "peft_gemm_bf16: null operand");
  PEFT_REQUIRE(K % 8 == 0, "peft_gemm_bf16: K must be a multiple of 8 (got %d)", K);
  PEFT_REQUIRE(lda >= K && ldb >= K, "peft_gemm_bf16: pitch smaller than K");
  const bool f32out = (epilogue == EPI_RESID_F32 || epilogue == EPI_STORE_F32);
  const int oelt = f32out ? 4 : 2;
  PEFT_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0 && (ldo * oelt) % 16 == 0,
               "peft_gemm_bf16: out must be 16-byte aligned with a 16-byte-multiple pitch");
  const bool needs_out2 = (epilogue == EPI_GELU || epilogue == EPI_QUICKGELU || epilogue == EPI_GELU_TANH);
  if (needs_out2)
    PEFT_REQUIRE(out2 && (reinterpret_cast<uintptr_t>(out2) & 15) == 0 && (ldo2 * 2) % 16 == 0,
                 "peft_gemm_bf16: activation epilogue needs an aligned out2");
  const bool needs_aux = (epilogue == EPI_RESID_F32 || epilogue == EPI_DGELU || epilogue == EPI_DQUICKGELU ||
                          epilogue == EPI_DRELU || epilogue == EPI_DGELU_TANH || epilogue == EPI_RESID_BF16);
  if (needs_aux) {
    const int aelt = (epilogue == EPI_RESID_F32) ? 4 : 2;
    PEFT_REQUIRE(aux && (reinterpret_cast<uintptr_t>(aux) & 15) == 0 && (ldaux * aelt) % 16 == 0,
                 "peft_gemm_bf16: epilogue %d needs an aligned aux operand", epilogue);
  }
  if (bias) PEFT_REQUIRE((reinterpret_cast<uintptr_t>(bias) & 15) == 0, "peft_gemm_bf16: bias must be 16-byte aligned");
  if (rowscale)
    PEFT_REQUIRE(epilogue == EPI_RESID_F32 || epilogue == EPI_STORE || epilogue == EPI_DGELU || epilogue == EPI_DQUICKGELU,
                 "peft_gemm_bf16_rs: rowscale is supported with the STORE / RESID_F32 / DGELU / DQUICKGELU epilogues (got %d)", epilogue);

  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  GemmParams p;
  p.M = M; p.N = N; p.K = K; p.bias = bias; p.out = out; p.ldo = ldo; p.out2 = out2; p.ldo2 = ldo2;
  p.aux = aux; p.ldaux = ldaux;
  p.rowscale = rowscale;
  p.ksplit = rowscale ? 1 : pick_ksplit(M, N, K, epilogue, bias, ldo);
  {
    static const int dbg = [] { const char* e = getenv("PEFT_GEMM_DBG"); return e ? atoi(e) : 0; }();
    p.dbg = dbg;
  }
  p.stats = gemm_stats_buffer();
  if (p.ksplit > 1) PEFT_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)M * N * sizeof(float), s));
  if (p.ksplit > 1 || use_gemm2(M, N)) {
    (void)sk_workspace();                     // allocated on the first 2-CTA launch of the device (outside any capture)
    decide_streamk(p, M, N, K);
    CUtensorMap ta, tb, to, to2, tx;
    int rc = make_tmap_2d(&ta, a, false, (unsigned long long)M, (unsigned long long)K, (unsigned long long)lda, kBM, 64);
    if (rc) return rc;
    rc = make_tmap_2d(&tb, b, false, (unsigned long long)N, (unsigned long long)K, (unsigned long long)ldb, kG2BN / 2, 64);
    if (rc) return rc;
    rc = make_tmap_2d(&to, out, f32out, (unsigned long long)M, (unsigned long long)N, (unsigned long long)ldo, 32, f32out ? 32 : 64);
    if (rc) return rc;
    to2 = to; tx = to;
    if (needs_out2) {
      rc = make_tmap_2d(&to2, out2, false, (unsigned long long)M, (unsigned long long)N, (unsigned long long)ldo2, 32, 64);
      if (rc) return rc;
    }
    if (needs_aux) {
      const bool af32 = (epilogue == EPI_RESID_F32);
      rc = make_tmap_2d(&tx, aux, af32, (unsigned long long)M, (unsigned long long)N, (unsigned long long)ldaux, 32, af32 ? 32 : 64);
      if (rc) return rc;
    }
    return dispatch_gemm2(epilogue, ta, tb, to, to2, tx, p, s);
  }
  const int bn = pick_bn(M, N);
  CUtensorMap ta, tb;
  int rc = make_tmap_bf16_2d(&ta, a, (unsigned long long)M, (unsigned long long)K, (unsigned long long)lda, kBM);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&tb, b, (unsigned long long)N, (unsigned long long)K, (unsigned long long)ldb, (unsigned)bn);
  if (rc) return rc;
  switch (bn) {
    case 256: return dispatch_epi<256>(epilogue, ta, tb, p, s);
    case 128: return dispatch_epi<128>(epilogue, ta, tb, p, s);
    default: return dispatch_epi<32>(epilogue, ta, tb, p, s);
  }
}

extern "C" int peft_set_gemm_streamk(int mode) {
  PEFT_REQUIRE(mode >= 0 && mode <= 2, "peft_set_gemm_streamk: mode must be 0 (off), 1 (automatic) or 2 (whenever possible)");
  streamk_mode().store(mode);
  return 0;
}

extern "C" int peft_set_pdl(int enabled) {
  peft::pdl_mode().store(enabled ? (int)peft::PDL_ALL : 0);
  return 0;
}

extern "C" int peft_set_gemm_algo(int algo) {
  PEFT_REQUIRE(algo >= 0 && algo <= 2, "peft_set_gemm_algo: algo must be 0 (auto), 1 (1-CTA) or 2 (2-CTA)");
  gemm_algo().store(algo);
  return 0;
}

extern "C" int peft_gemm_tn_bf16(const void* x, long long ldx, const void* y, long long ldy, int M, int N, int T, float* out,
                                 long long ldo, void* stream) {
  PEFT_REQUIRE(M > 0 && N > 0 && T > 0 && x && y && out, "peft_gemm_tn_bf16: bad shape M=%d N=%d T=%d or null operand", M, N, T);
  PEFT_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0 && ldx % 8 == 0 &&
                   ldy % 8 == 0 && ldx >= M && ldy >= N,
               "peft_gemm_tn_bf16: operands must be 16-byte aligned with pitches %% 8 == 0");
  PEFT_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0 && ldo == N && N % 4 == 0,
               "peft_gemm_tn_bf16: out must be contiguous [M, N], 16-byte aligned, N %% 4 == 0");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  GemmParams p{};
  p.M = M; p.N = N; p.K = T; p.bias = nullptr; p.out = out; p.ldo = ldo;
  p.ksplit = pick_ksplit(M, N, T, EPI_STORE_F32, nullptr, ldo);
  if (p.ksplit > 1) PEFT_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)M * N * sizeof(float), s));
  CUtensorMap ta, tb, to;
  int rc = make_tmap_2d(&ta, x, false, (unsigned long long)T, (unsigned long long)M, (unsigned long long)ldx, 64, 64);
  if (rc) return rc;
  rc = make_tmap_2d(&tb, y, false, (unsigned long long)T, (unsigned long long)N, (unsigned long long)ldy, 64, 64);
  if (rc) return rc;
  rc = make_tmap_2d(&to, out, true, (unsigned long long)M, (unsigned long long)N, (unsigned long long)ldo, 32, 32);
  if (rc) return rc;
  return launch_gemm2<EPI_STORE_F32, true>(ta, tb, to, to, to, p, s);
}

extern "C" int peft_debug_gemm_stats(unsigned long long* host_out, int n_ctas) {
#if !PEFT_GEMM_DIAG
  return peft::fail(PEFT_ERR_INVALID, "peft_debug_gemm_stats: the library was built without the in-kernel GEMM diagnostics (make DIAG=1)");
#endif
  unsigned long long* buf = peft::gemm_stats_buffer();
  PEFT_REQUIRE(buf != nullptr, "peft_debug_gemm_stats: set PEFT_GEMM_STATS=1 before the first GEMM");
  PEFT_REQUIRE(host_out && n_ctas > 0 && n_ctas <= 256, "peft_debug_gemm_stats: bad arguments");
  PEFT_CUDA_CHECK(cudaDeviceSynchronize());
  PEFT_CUDA_CHECK(cudaMemcpy(host_out, buf, (size_t)n_ctas * 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return 0;
}

namespace peft {
// ---------------------------------------------------------------- allocation-free stream-K product
// peft_gemm_bf16_noalloc: the 2-CTA product of peft_gemm_bf16 for callers that may not allocate device memory (code under
// CUDA-graph capture that never ran a warm-up call, serving loops, harnesses that forbid cudaMalloc).  The stream-K fix-up
// scratch — one [256 x 256] fp32 partial-tile slot per SM pair + the per-warp flags — lives in the file-scope device arrays
// below instead of the cudaMalloc'ed per-device workspace of sk_workspace(): kNoAllocMaxClusters slots = 20 MiB + 5 KiB,
// held by every GPU that loads the library until the process exits.  No counterpart in the
// reference.
//   * In host code `g_noalloc_ws` is only a stand-in for the symbol: its device address is taken once per device, with that
//     device current, through cudaGetSymbolAddress (noalloc_scratch below) and handed to the kernel in GemmParams.
//   * All calls on a device share the one array: two calls whose kernels overlap on the GPU (different streams, or graph
//     branches) race.  Serialise them on one stream.  The array keeps the last call's partial sums — every slot is fully
//     written before its flag is raised, so nothing but the flags must start at zero; they are cleared on the stream before
//     every launch (a launch that was aborted by the watchdog trap cannot poison the next one).
constexpr int kNoAllocMaxClusters = 80;                                       // B200: 148 SMs = 74 pairs
constexpr size_t kNoAllocSlotF4 = (size_t)2 * kEpiWarps * 1024;               // float4 per cluster slot (256 KiB)
static __device__ __align__(256) float4 g_noalloc_ws[kNoAllocMaxClusters * kNoAllocSlotF4];
static __device__ __align__(256) int g_noalloc_flags[kNoAllocMaxClusters * 2 * kEpiWarps];

struct NoAllocScratch { float4* ws = nullptr; int* flags = nullptr; };
static int noalloc_scratch(NoAllocScratch* out) {
  static std::mutex mu;
  static NoAllocScratch per_dev[64];
  int dev = 0;
  PEFT_CUDA_CHECK(cudaGetDevice(&dev));
  PEFT_REQUIRE(dev >= 0 && dev < 64, "peft_gemm_bf16_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> g(mu);
  NoAllocScratch& s = per_dev[dev];
  if (s.ws == nullptr) {                      // once per device, with that device current
    void* pw = nullptr;
    void* pf = nullptr;
    PEFT_CUDA_CHECK(cudaGetSymbolAddress(&pw, g_noalloc_ws));
    PEFT_CUDA_CHECK(cudaGetSymbolAddress(&pf, g_noalloc_flags));
    PEFT_REQUIRE(pw != nullptr && pf != nullptr, "peft_gemm_bf16_noalloc: cudaGetSymbolAddress returned a null address");
    s.flags = reinterpret_cast<int*>(pf);
    s.ws = reinterpret_cast<float4*>(pw);
  }
  *out = s;
  return 0;
}

}  // namespace peft

extern "C" int peft_gemm_bf16_noalloc(const void* a, long long lda, const void* b, long long ldb, int M, int N, int K,
                                      const float* bias, void* out, long long ldo, void* out2, long long ldo2,
                                      const void* aux, long long ldaux, int epilogue, void* stream) {
  PEFT_REQUIRE(M > 0 && N > 0 && K > 0, "peft_gemm_bf16_noalloc: empty problem M=%d N=%d K=%d", M, N, K);
  PEFT_REQUIRE(a && b && out, "peft_gemm_bf16_noalloc: null operand");
  PEFT_REQUIRE(epilogue >= 0 && epilogue <= EPI_RESID_BF16, "peft_gemm_bf16_noalloc: unknown epilogue %d", epilogue);
  PEFT_REQUIRE(K % 8 == 0, "peft_gemm_bf16_noalloc: K must be a multiple of 8 (got %d)", K);
  PEFT_REQUIRE(lda >= K && ldb >= K, "peft_gemm_bf16_noalloc: pitch smaller than K");
  const int clusters = sm_count() / 2;
  PEFT_REQUIRE(clusters <= kNoAllocMaxClusters,
               "peft_gemm_bf16_noalloc: this device has %d SM pairs, the static stream-K scratch holds %d slots (20 MiB); use "
               "peft_gemm_bf16", clusters, kNoAllocMaxClusters);
  const long long tiles = (long long)((M + 255) / 256) * ((N + 255) / 256);
  PEFT_REQUIRE(tiles <= (1 << 24), "peft_gemm_bf16_noalloc: %lld output tiles exceed the supported maximum of %d", tiles, 1 << 24);
  const bool f32out = (epilogue == EPI_RESID_F32 || epilogue == EPI_STORE_F32);
  const int oelt = f32out ? 4 : 2;
  PEFT_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0 && (ldo * oelt) % 16 == 0,
               "peft_gemm_bf16_noalloc: out must be 16-byte aligned with a 16-byte-multiple pitch");
  const bool needs_out2 = (epilogue == EPI_GELU || epilogue == EPI_QUICKGELU || epilogue == EPI_GELU_TANH);
  if (needs_out2)
    PEFT_REQUIRE(out2 && (reinterpret_cast<uintptr_t>(out2) & 15) == 0 && (ldo2 * 2) % 16 == 0,
                 "peft_gemm_bf16_noalloc: activation epilogue needs an aligned out2");
  const bool needs_aux = (epilogue == EPI_RESID_F32 || epilogue == EPI_DGELU || epilogue == EPI_DQUICKGELU ||
                          epilogue == EPI_DRELU || epilogue == EPI_DGELU_TANH || epilogue == EPI_RESID_BF16);
  if (needs_aux) {
    const int aelt = (epilogue == EPI_RESID_F32) ? 4 : 2;
    PEFT_REQUIRE(aux && (reinterpret_cast<uintptr_t>(aux) & 15) == 0 && (ldaux * aelt) % 16 == 0,
                 "peft_gemm_bf16_noalloc: epilogue %d needs an aligned aux operand", epilogue);
  }
  if (bias) PEFT_REQUIRE((reinterpret_cast<uintptr_t>(bias) & 15) == 0, "peft_gemm_bf16_noalloc: bias must be 16-byte aligned");

  NoAllocScratch sc;
  int rc = noalloc_scratch(&sc);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  GemmParams p;
  p.M = M; p.N = N; p.K = K; p.bias = bias; p.out = out; p.ldo = ldo; p.out2 = out2; p.ldo2 = ldo2;
  p.aux = aux; p.ldaux = ldaux;
  p.ksplit = 1;
  // stream-K whenever the shape allows it (the rule of decide_streamk's mode 2): all but the last full wave stay data-parallel,
  // the k-blocks of the rest are dealt out evenly; otherwise a plain data-parallel launch that touches no scratch at all
  p.sk_dp_tiles = -1;
  const int KB = (K + kBK - 1) / kBK;
  const long long dp = tiles >= clusters ? (tiles / clusters - 1) * clusters : 0;
  if (KB >= 2 && (tiles - dp) * KB >= 4LL * clusters) {
    p.sk_dp_tiles = (int)dp;
    p.sk_ws = sc.ws;
    p.sk_flags = sc.flags;
    PEFT_CUDA_CHECK(cudaMemsetAsync(sc.flags, 0, sizeof(int) * kNoAllocMaxClusters * 2 * kEpiWarps, s));
  }
  CUtensorMap ta, tb, to, to2, tx;
  rc = make_tmap_2d(&ta, a, false, (unsigned long long)M, (unsigned long long)K, (unsigned long long)lda, kBM, 64);
  if (rc) return rc;
  rc = make_tmap_2d(&tb, b, false, (unsigned long long)N, (unsigned long long)K, (unsigned long long)ldb, kG2BN / 2, 64);
  if (rc) return rc;
  rc = make_tmap_2d(&to, out, f32out, (unsigned long long)M, (unsigned long long)N, (unsigned long long)ldo, 32, f32out ? 32 : 64);
  if (rc) return rc;
  to2 = to; tx = to;
  if (needs_out2) {
    rc = make_tmap_2d(&to2, out2, false, (unsigned long long)M, (unsigned long long)N, (unsigned long long)ldo2, 32, 64);
    if (rc) return rc;
  }
  if (needs_aux) {
    const bool af32 = (epilogue == EPI_RESID_F32);
    rc = make_tmap_2d(&tx, aux, af32, (unsigned long long)M, (unsigned long long)N, (unsigned long long)ldaux, 32, af32 ? 32 : 64);
    if (rc) return rc;
  }
  return dispatch_gemm2(epilogue, ta, tb, to, to2, tx, p, s);
}
