// C-ABI: CLIP-dialect delta placement and the Fastfood projection.
#include <stdlib.h>
#include "common.cuh"
#include "clip_delta_sm100.cuh"
#include "fastfood_sm100.cuh"
#include "allreduce_sgd_sm100.cuh"
#include "preprocess_sm100.cuh"
#include "factors_sm100.cuh"

using namespace peft;

namespace {
inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// persistent grid: occupancy x SMs CTAs, launched cooperatively so that every CTA is resident (the in-kernel waits rely on it)
template <typename K>
int ff_launch(K kernel, FfArgs a, cudaStream_t s) {
  static DeviceMax dev_smem;
  const int smem = ff_smem_bytes(a.wn);
  if (dev_smem.get() < smem) {   // raise the opt-in limit once per device
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ff_smem_bytes(kFfMaxWn)));
    dev_smem.set(ff_smem_bytes(kFfMaxWn));
  }
  int occ = 0;
  PEFT_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kFfThreads, smem));
  PEFT_REQUIRE(occ >= 1, "fastfood: kernel does not fit on an SM (smem %d B)", smem);
  long long grid = (long long)occ * sm_count();
  if (grid > a.n_items) grid = a.n_items;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kFfThreads);
  cfg.dynamicSmemBytes = (size_t)smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  PEFT_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kernel, a));
  PEFT_LAUNCH_CHECK();
  return 0;
}
int ff_discard() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("PEFT_FF_DISCARD");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v;
}
int ff_check(int d, int wn, int T, int n_items) {
  PEFT_REQUIRE(d > 0 && wn >= d && wn <= kFfMaxWn && (wn & (wn - 1)) == 0, "fastfood: need d <= wn <= %d, wn a power of two (d=%d wn=%d)", kFfMaxWn, d, wn);
  PEFT_REQUIRE(T > 0 && n_items > 0, "fastfood: empty tensor / item list");
  return 0;
}
}  // namespace

extern "C" int peft_clip_delta(void* qkv, long long ldqkv, void* delta, long long ldd, int B, int H, int L, float mult_q,
                               float mult_v, int backward, void* stream) {
  PEFT_REQUIRE(B > 0 && H > 0 && L > 0, "peft_clip_delta: bad shape B=%d H=%d L=%d", B, H, L);
  PEFT_REQUIRE(qkv && delta && al16(qkv) && al16(delta) && ldqkv % 8 == 0 && ldd % 8 == 0 && ldqkv >= 3ll * H * 64 && ldd >= 2ll * H * 64,
               "peft_clip_delta: operands must be 16-byte aligned, pitch %% 8 == 0, qkv >= 3*H*64 and delta >= 2*H*64 columns");
  const long long total = (long long)B * H * L * 16;
  long long blocks = (total + 255) / 256;
  if (blocks > sm_count() * 16) blocks = sm_count() * 16;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (backward)
    (void)launch_pdl(PDL_MISC, clip_delta_kernel<true>, dim3((int)blocks), dim3(256), 0, s, (__nv_bfloat16*)qkv, ldqkv, (__nv_bfloat16*)delta, ldd, B, H, L, mult_q, mult_v);
  else
    (void)launch_pdl(PDL_MISC, clip_delta_kernel<false>, dim3((int)blocks), dim3(256), 0, s, (__nv_bfloat16*)qkv, ldqkv, (__nv_bfloat16*)delta, ldd, B, H, L, mult_q, mult_v);
  PEFT_LAUNCH_CHECK();
  return 0;
}

static int transpose_impl(const void* in, long long ldin, int rows, int cols, void* out, long long ldout, float* colsum,
                          void* stream) {
  PEFT_REQUIRE(rows > 0 && cols > 0 && in && out, "peft_transpose_bf16: bad shape rows=%d cols=%d", rows, cols);
  PEFT_REQUIRE(ldin % 2 == 0 && ldout % 2 == 0 && ldin >= cols && ldout >= rows && (reinterpret_cast<uintptr_t>(in) & 3) == 0 &&
                   (reinterpret_cast<uintptr_t>(out) & 3) == 0,
               "peft_transpose_bf16: pitches must be even and pointers 4-byte aligned");
  const bool vec = al16(in) && al16(out) && ldin % 8 == 0 && ldout % 8 == 0;
  PEFT_REQUIRE(colsum == nullptr || vec, "peft_transpose_colsum_bf16: needs 16-byte aligned operands and pitches %% 8 == 0");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (vec) {
    dim3 grid((cols + 255) / 256, (rows + 63) / 64);
    if (colsum)
      (void)launch_pdl(PDL_MISC, transpose_bf16_vec_kernel<true>, dim3(grid), dim3(256), 0, s, (const __nv_bfloat16*)in, ldin, rows, cols, (__nv_bfloat16*)out, ldout, colsum);
    else
      (void)launch_pdl(PDL_MISC, transpose_bf16_vec_kernel<false>, dim3(grid), dim3(256), 0, s, (const __nv_bfloat16*)in, ldin, rows, cols, (__nv_bfloat16*)out, ldout, nullptr);
  } else {
    dim3 grid((cols + 63) / 64, (rows + 63) / 64);
    (void)launch_pdl(PDL_MISC, transpose_bf16_kernel, dim3(grid), dim3(256), 0, s, (const __nv_bfloat16*)in, ldin, rows, cols, (__nv_bfloat16*)out, ldout);
  }
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_transpose_bf16(const void* in, long long ldin, int rows, int cols, void* out, long long ldout, void* stream) {
  return transpose_impl(in, ldin, rows, cols, out, ldout, nullptr, stream);
}

extern "C" int peft_transpose_colsum_bf16(const void* in, long long ldin, int rows, int cols, void* out, long long ldout,
                                          float* colsum, void* stream) {
  PEFT_REQUIRE(colsum != nullptr, "peft_transpose_colsum_bf16: null colsum");
  return transpose_impl(in, ldin, rows, cols, out, ldout, colsum, stream);
}

extern "C" int peft_colsum_bf16(const void* in, long long ldin, int rows, int cols, float* colsum, void* stream) {
  PEFT_REQUIRE(rows > 0 && cols > 0 && in && colsum, "peft_colsum_bf16: bad shape rows=%d cols=%d or null operand", rows, cols);
  PEFT_REQUIRE(al16(in) && ldin % 8 == 0 && ldin >= cols, "peft_colsum_bf16: needs a 16-byte aligned operand and pitch %% 8 == 0");
  const int colblocks = (cols + 255) / 256, nrb = (rows + 63) / 64;
  // ~4 CTAs per SM in total, each walking nrb / gy row blocks, but at most 100 CTAs (= same-address atomics) per column block.
  // Measured r02 at 12 800 rows (graph replay): 768 columns 10.9 us (2 / SM, gy = 99) vs 17.6 us (4 / SM, gy = 198); 2304 columns
  // 13.8 -> 10.7 us and 3072 columns 19.7 -> 12.4 us going from 2 to 4 CTAs per SM; 8 and 16 per SM are slower again.
  static const int per_sm = [] { const char* e = getenv("PEFT_COLSUM_CTAS"); return (e && atoi(e) > 0) ? atoi(e) : 4; }();
  int gy = (per_sm * sm_count() + colblocks - 1) / colblocks;
  if (gy > 100) gy = 100;
  if (gy > nrb) gy = nrb;
  if (gy < 1) gy = 1;
  (void)launch_pdl(PDL_MISC, colsum_bf16_kernel, dim3(dim3(colblocks, gy)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), (const __nv_bfloat16*)in, ldin, rows,
                                                                                             cols, colsum);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_phm_fold(const float* rule, const float* s_or_dH, float* out, int n, int backward, void* stream) {
  PEFT_REQUIRE(n > 0 && n <= kFoldMaxN && ((long long)n * n) % 4 == 0, "peft_phm_fold: need n <= %d and n*n %% 4 == 0 (n=%d)", kFoldMaxN, n);
  PEFT_REQUIRE(rule && s_or_dH && out && al16(rule) && al16(s_or_dH) && al16(out), "peft_phm_fold: null or unaligned pointer");
  const long long nn = (long long)n * n;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!backward) {
    const long long threads = nn / 4;
    phm_fold_fwd_kernel<<<(unsigned)((threads + kFoldThreads - 1) / kFoldThreads), kFoldThreads, 0, st>>>(rule, s_or_dH, out, n, nn);
  } else {
    PEFT_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)2 * n * sizeof(float), st));
    phm_fold_bwd_kernel<<<dim3((n + kFoldRows - 1) / kFoldRows, kFoldSplit), kFoldThreads, 0, st>>>(rule, s_or_dH, out, n, nn);
  }
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_fastfood_fwd(const float* V, int d, int wn, const peft_ff_tensor* tensors, int T, const int* items,
                                 int n_items, const uint16_t* pi_lo, const float* gg, const float* bsign,
                                 const float* theta0, float* theta, float* wu, float* scratch, int* counters, void* stream) {
  int rc = ff_check(d, wn, T, n_items);
  if (rc) return rc;
  PEFT_REQUIRE(V && tensors && items && pi_lo && gg && bsign && theta0 && theta && wu && scratch && counters, "peft_fastfood_fwd: null operand");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  PEFT_CUDA_CHECK(cudaMemsetAsync(counters, 0, (size_t)T * sizeof(int), s));
  ff_small_kernel<0><<<T, 256, wn * 4, s>>>(tensors, bsign, V, wu, nullptr, d, wn);
  PEFT_LAUNCH_CHECK();
  FfArgs a{};
  a.tensors = tensors; a.items = reinterpret_cast<const int2*>(items); a.pi_lo = pi_lo; a.gg = gg; a.wv = wu; a.scratch = scratch;
  a.theta0 = theta0; a.theta = theta; a.counters = counters; a.wn = wn; a.T = T; a.n_items = n_items; a.discard = ff_discard();
  return ff_launch(ff_fwd_kernel, a, s);
}

extern "C" int peft_fastfood_bwd(const float* dtheta, int d, int wn, const peft_ff_tensor* tensors, int T, const int* items,
                                 int n_items, const uint16_t* pi_lo, const float* gg, const float* bsign, float* wu,
                                 float* scratch, int* counters, float* dV, void* stream) {
  int rc = ff_check(d, wn, T, n_items);
  if (rc) return rc;
  PEFT_REQUIRE(dtheta && tensors && items && pi_lo && gg && bsign && wu && scratch && counters && dV, "peft_fastfood_bwd: null operand");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  PEFT_CUDA_CHECK(cudaMemsetAsync(counters, 0, (size_t)(T + 1) * sizeof(int), s));     // [T] per-tensor counters + the item ticket
  PEFT_CUDA_CHECK(cudaMemsetAsync(wu, 0, (size_t)T * wn * 4, s));
  FfArgs a{};
  a.tensors = tensors; a.items = reinterpret_cast<const int2*>(items); a.pi_lo = pi_lo; a.gg = gg; a.u = wu; a.scratch = scratch;
  a.dtheta = dtheta; a.counters = counters; a.wn = wn; a.T = T; a.n_items = n_items;
  static DeviceOnce bwd_attr;
  if (bwd_attr.needs()) {
    PEFT_CUDA_CHECK(cudaFuncSetAttribute(ff_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (kFfTile + kFfMaxWn) * 4));
    bwd_attr.done();
  }
  ff_bwd_kernel<<<n_items, kFfThreads, (kFfTile + wn) * 4, s>>>(a);
  PEFT_LAUNCH_CHECK();
  ff_small_kernel<1><<<T, 256, wn * 4, s>>>(tensors, bsign, nullptr, wu, dV, d, wn);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_allreduce_sgd(const void* const* peer_grads, void* const* peer_pads, int rank, int world, float* grad,
                                  float* param, float* momentum, long long n, const float* lr, float mu, float weight_decay,
                                  int zero_grad, int pad_base, int pad_slots, void* stream) {
  PEFT_REQUIRE(world >= 1 && world <= kArMaxWorld && rank >= 0 && rank < world, "peft_allreduce_sgd: bad rank %d / world %d", rank, world);
  PEFT_REQUIRE(grad && param && momentum && lr && n > 0 && n % 4 == 0, "peft_allreduce_sgd: null operand or n %% 4 != 0 (n=%lld)", n);
  PEFT_REQUIRE(al16(grad) && al16(param) && al16(momentum), "peft_allreduce_sgd: flat buffers must be 16-byte aligned");
  PEFT_REQUIRE(world == 1 || (peer_grads && peer_pads), "peft_allreduce_sgd: peer pointer tables missing");
  long long blocks = (n / 4 + 4 * kArThreads - 1) / (4 * kArThreads);      // >= 4 float4 per thread
  if (blocks < 1) blocks = 1;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  PEFT_REQUIRE(world == 1 || pad_base + 2 * blocks * world <= pad_slots, "peft_allreduce_sgd: signal pad too small (%d slots, need %lld)",
               pad_slots, pad_base + 2 * blocks * world);
  AllreduceSgdArgs a{};
  a.peer_grads = reinterpret_cast<const float* const*>(peer_grads);
  a.peer_pads = reinterpret_cast<uint32_t* const*>(peer_pads);
  a.grad = grad; a.param = param; a.momentum = momentum; a.lr = lr; a.n = n; a.rank = rank; a.world = world;
  a.mu = mu; a.weight_decay = weight_decay; a.zero_grad = zero_grad; a.pad_base = pad_base;
  allreduce_sgd_kernel<<<(unsigned)blocks, kArThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_preprocess_clip(const uint8_t* img, int H, int W, const int* hb, const int* hk, int hksize, const int* vb,
                                    const int* vk, int vksize, int top, int left, int n, int r0, int r1, const float* mean,
                                    const float* stdv, uint8_t* tmp, float* out, void* stream) {
  PEFT_REQUIRE(img && hb && hk && vb && vk && mean && stdv && tmp && out, "peft_preprocess_clip: null operand");
  PEFT_REQUIRE(H > 0 && W > 0 && n > 0 && hksize > 0 && vksize > 0 && top >= 0 && left >= 0 && r0 >= 0 && r1 > r0 && r1 <= H,
               "peft_preprocess_clip: bad geometry H=%d W=%d n=%d rows [%d, %d)", H, W, n, r0, r1);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long t1 = (long long)(r1 - r0) * n * 3, t2 = (long long)n * n * 3;
  const int cap = sm_count() * 8;
  long long b1 = (t1 + 255) / 256, b2 = (t2 + 255) / 256;
  if (b1 > cap) b1 = cap;
  if (b2 > cap) b2 = cap;
  pp_horizontal_kernel<<<(unsigned)b1, 256, 0, s>>>(img, W, hb, hk, hksize, left, n, r0, r1, tmp);
  PEFT_LAUNCH_CHECK();
  pp_vertical_kernel<<<(unsigned)b2, 256, 0, s>>>(tmp, vb, vk, vksize, top, n, r0, mean, stdv, out);
  PEFT_LAUNCH_CHECK();
  return 0;
}


// Reshape-trick operands of the small-n KAdaptation update for all layers (factors_sm100.cuh).  backward == 0: a = wl1, b = wr1,
// c = wl2, d = wr2 (inputs), out0 = Dn, out1 = UpT;  backward != 0: out0 = dDn, out1 = dUpT (inputs), a..d = the four gradients.
extern "C" int peft_kron_factors(const float* rule, float* a, float* b, float* c, float* d, float* out0, float* out1, int dep, int n,
                                 int K, int r, int P, int backward, void* stream) {
  PEFT_REQUIRE(dep > 0 && n > 0 && K > 0 && r > 0 && P > 0, "peft_kron_factors: bad shape dep=%d n=%d K=%d r=%d P=%d", dep, n, K, r, P);
  PEFT_REQUIRE(rule && a && b && c && d && out0 && out1, "peft_kron_factors: null operand");
  PEFT_REQUIRE(dep <= 65535 && 2ll * n * r * n <= 2147483647ll, "peft_kron_factors: too many layers / rows");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (!backward) {
    kron_factors_fwd_kernel<<<dim3(2 * n * r * n, dep), 256, 0, s>>>(rule, a, b, c, d, out0, out1, n, K, r, P);
  } else {
    const long long total = 2ll * dep * n * K * r + 2ll * dep * n * r * P;
    kron_factors_bwd_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(rule, out0, out1, a, b, c, d, dep, n, K, r, P);
  }
  PEFT_LAUNCH_CHECK();
  return 0;
}

// gradient of a trainable phm_rule [dep,n,n,n] from dDn and the two left factors (written, not accumulated)
extern "C" int peft_kron_factors_drule(const float* dDn, const float* wl1, const float* wl2, float* drule, int dep, int n, int K, int r,
                                       void* stream) {
  PEFT_REQUIRE(dep > 0 && n > 0 && K > 0 && r > 0 && dDn && wl1 && wl2 && drule, "peft_kron_factors_drule: bad shape or null operand");
  const long long warps = (long long)dep * n * n * n;
  kron_factors_drule_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(dDn, wl1, wl2, drule,
                                                                                                                    dep, n, K, r);
  PEFT_LAUNCH_CHECK();
  return 0;
}
