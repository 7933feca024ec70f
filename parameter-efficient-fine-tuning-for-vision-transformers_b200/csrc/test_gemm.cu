// Standalone GPU self-test of peft_gemm_bf16 against a CPU double-precision reference (sampled entries).
// Not part of the product; used to validate the tcgen05/TMA kernel early and to time it in isolation.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../include/peft_b200.h"

static float bf16r(float x) { return __bfloat162float(__float2bfloat16(x)); }
static double gelu(double x) { return 0.5 * x * (1.0 + erf(x / sqrt(2.0))); }
static double qgelu(double x) { return x / (1.0 + exp(-1.702 * x)); }
static double dqgelu(double x) { double s = 1.0 / (1.0 + exp(-1.702 * x)); return s * (1.0 + 1.702 * x * (1.0 - s)); }
static double tgelu(double x) { return 0.5 * x * (1.0 + tanh(sqrt(2.0 / M_PI) * (x + 0.044715 * x * x * x))); }
static double dtgelu(double x) { double u = sqrt(2.0 / M_PI) * (x + 0.044715 * x * x * x), t = tanh(u); return 0.5 * (1 + t) + 0.5 * x * (1 - t * t) * sqrt(2.0 / M_PI) * (1 + 3 * 0.044715 * x * x); }
static double dgelu(double x) { return 0.5 * (1.0 + erf(x / sqrt(2.0))) + x * exp(-0.5 * x * x) / sqrt(2.0 * M_PI); }

struct Case { int M, N, K, epi; int lda_pad; };

int run_case(const Case& c, bool timeit) {
  const int M = c.M, N = c.N, K = c.K;
  const long long lda = K + c.lda_pad, ldb = K + c.lda_pad;
  std::vector<__nv_bfloat16> hA((size_t)M * lda), hB((size_t)N * ldb), hAux((size_t)M * N);
  std::vector<float> hbias(N), hres((size_t)M * N);
  srand(1234 + M + N + K);
  auto rnd = []() { return (rand() / (float)RAND_MAX) * 2.f - 1.f; };
  for (auto& v : hA) v = __float2bfloat16(rnd());
  for (auto& v : hB) v = __float2bfloat16(rnd() * 0.1f);
  for (auto& v : hbias) v = rnd();
  for (auto& v : hAux) v = __float2bfloat16(rnd() * 2.f);
  for (auto& v : hres) v = rnd();
  __nv_bfloat16 *dA, *dB, *dAux, *dO, *dO2; float *dbias, *dres, *dOf;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dAux, hAux.size() * 2);
  cudaMalloc(&dO, (size_t)M * N * 2); cudaMalloc(&dO2, (size_t)M * N * 2); cudaMalloc(&dbias, N * 4);
  cudaMalloc(&dres, (size_t)M * N * 4); cudaMalloc(&dOf, (size_t)M * N * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dAux, hAux.data(), hAux.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dbias, hbias.data(), N * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dres, hres.data(), (size_t)M * N * 4, cudaMemcpyHostToDevice);
  cudaMemset(dO, 0, (size_t)M * N * 2); cudaMemset(dO2, 0, (size_t)M * N * 2); cudaMemset(dOf, 0, (size_t)M * N * 4);
  const bool f32 = (c.epi == PEFT_EPI_RESID_F32 || c.epi == PEFT_EPI_STORE_F32);
  const void* aux = (c.epi == PEFT_EPI_RESID_F32) ? (const void*)dres : (const void*)dAux;
  void* out = f32 ? (void*)dOf : (void*)dO;
  int rc = peft_gemm_bf16(dA, lda, dB, ldb, M, N, K, dbias, out, N, dO2, N, aux, N, c.epi, nullptr);
  if (rc) { char e[512]; peft_last_error(e, 512); printf("FAIL rc=%d %s\n", rc, e); return 1; }
  cudaError_t ce = cudaDeviceSynchronize();
  if (ce != cudaSuccess) { printf("FAIL cuda: %s\n", cudaGetErrorString(ce)); return 1; }
  std::vector<__nv_bfloat16> hO((size_t)M * N), hO2((size_t)M * N); std::vector<float> hOf((size_t)M * N);
  cudaMemcpy(hO.data(), dO, hO.size() * 2, cudaMemcpyDeviceToHost);
  cudaMemcpy(hO2.data(), dO2, hO2.size() * 2, cudaMemcpyDeviceToHost);
  cudaMemcpy(hOf.data(), dOf, hOf.size() * 4, cudaMemcpyDeviceToHost);
  double max_err = 0; int bad = 0;
  const int samples = 4000;
  for (int s = 0; s < samples; ++s) {
    int i, j;
    if (s < 64) { i = (s & 1) ? M - 1 - (s / 8) % M : (s / 8) % M; j = (s & 2) ? N - 1 - (s % 8) % N : (s % 8) % N; if (i < 0) i = 0; if (j < 0) j = 0; }
    else { i = rand() % M; j = rand() % N; }
    double acc = 0;
    for (int k = 0; k < K; ++k) acc += (double)__bfloat162float(hA[(size_t)i * lda + k]) * (double)__bfloat162float(hB[(size_t)j * ldb + k]);
    acc += hbias[j];
    double ref = acc, got, ref2 = 0, got2 = 0; bool two = false;
    const double h = __bfloat162float(hAux[(size_t)i * N + j]);
    switch (c.epi) {
      case PEFT_EPI_STORE: got = __bfloat162float(hO[(size_t)i * N + j]); break;
      case PEFT_EPI_GELU: got = __bfloat162float(hO[(size_t)i * N + j]); two = true; ref2 = gelu(bf16r((float)acc)); got2 = __bfloat162float(hO2[(size_t)i * N + j]); break;
      case PEFT_EPI_RESID_F32: ref = acc + hres[(size_t)i * N + j]; got = hOf[(size_t)i * N + j]; break;
      case PEFT_EPI_STORE_F32: got = hOf[(size_t)i * N + j]; break;
      case PEFT_EPI_DGELU: ref = acc * dgelu(h); got = __bfloat162float(hO[(size_t)i * N + j]); break;
      case PEFT_EPI_RESID_BF16: ref = acc + h; got = __bfloat162float(hO[(size_t)i * N + j]); break;
      case PEFT_EPI_QUICKGELU: got = __bfloat162float(hO[(size_t)i * N + j]); two = true; ref2 = qgelu(bf16r((float)acc)); got2 = __bfloat162float(hO2[(size_t)i * N + j]); break;
      case PEFT_EPI_GELU_TANH: got = __bfloat162float(hO[(size_t)i * N + j]); two = true; ref2 = tgelu(bf16r((float)acc)); got2 = __bfloat162float(hO2[(size_t)i * N + j]); break;
      case PEFT_EPI_DQUICKGELU: ref = acc * dqgelu(h); got = __bfloat162float(hO[(size_t)i * N + j]); break;
      case PEFT_EPI_DGELU_TANH: ref = acc * dtgelu(h); got = __bfloat162float(hO[(size_t)i * N + j]); break;
      case PEFT_EPI_RELU: ref = acc > 0 ? acc : 0; got = __bfloat162float(hO[(size_t)i * N + j]); break;
      case PEFT_EPI_DRELU: ref = h > 0 ? acc : 0; got = __bfloat162float(hO[(size_t)i * N + j]); break;
      default: got = ref;
    }
    const double tol = f32 ? 2e-3 : 2e-2;
    double err = fabs(got - ref) / (fabs(ref) + 1.0);
    if (two) err = fmax(err, fabs(got2 - ref2) / (fabs(ref2) + 1.0));
    if (err > max_err) max_err = err;
    if (err > tol) { if (bad < 5) printf("  mismatch (%d,%d): got %f ref %f\n", i, j, got, ref); ++bad; }
  }
  printf("%s M=%d N=%d K=%d epi=%d pad=%d max_rel_err=%.3e bad=%d\n", bad ? "FAIL" : "PASS", M, N, K, c.epi, c.lda_pad, max_err, bad);
  if (timeit && !bad) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) peft_gemm_bf16(dA, lda, dB, ldb, M, N, K, dbias, out, N, dO2, N, aux, N, c.epi, nullptr);
    cudaEventRecord(e0);
    const int iters = 20;
    for (int i = 0; i < iters; ++i) peft_gemm_bf16(dA, lda, dB, ldb, M, N, K, dbias, out, N, dO2, N, aux, N, c.epi, nullptr);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= iters;
    printf("  time %.3f us  %.1f TFLOP/s\n", ms * 1e3, 2.0 * M * N * K / (ms * 1e-3) / 1e12);
  }
  cudaFree(dA); cudaFree(dB); cudaFree(dAux); cudaFree(dO); cudaFree(dO2); cudaFree(dbias); cudaFree(dres); cudaFree(dOf);
  return bad ? 1 : 0;
}

int main(int argc, char** argv) {
  const bool timeit = argc > 1;
  const int algo = argc > 2 ? atoi(argv[2]) : 0;
  if (peft_set_gemm_algo(algo)) { printf("bad algo\n"); return 2; }
  printf("== gemm self-test, algo=%d (0 auto, 1 = 1-CTA, 2 = 2-CTA)\n", algo);
  std::vector<Case> cases = {
      {128, 128, 64, PEFT_EPI_STORE, 0},   {128, 256, 128, PEFT_EPI_STORE, 0}, {256, 128, 768, PEFT_EPI_STORE, 0},
      {1000, 768, 768, PEFT_EPI_STORE, 0}, {1576, 2304, 800, PEFT_EPI_STORE, 32}, {1576, 32, 768, PEFT_EPI_STORE, 0},
      {1576, 16, 768, PEFT_EPI_STORE, 0},  {777, 200, 72, PEFT_EPI_STORE, 8},
      {1576, 3072, 768, PEFT_EPI_GELU, 0}, {1576, 768, 3072, PEFT_EPI_RESID_F32, 0}, {1576, 3072, 768, PEFT_EPI_DGELU, 0},
      {1576, 768, 768, PEFT_EPI_STORE_F32, 0}, {1576, 768, 768, PEFT_EPI_RESID_BF16, 0},
      {1576, 3072, 768, PEFT_EPI_QUICKGELU, 0}, {1576, 3072, 768, PEFT_EPI_DQUICKGELU, 0}, {1576, 3072, 768, PEFT_EPI_GELU_TANH, 0},
      {1576, 3072, 768, PEFT_EPI_DGELU_TANH, 0}, {1576, 1024, 768, PEFT_EPI_RELU, 0}, {1576, 1024, 768, PEFT_EPI_DRELU, 0},
      {6304, 768, 768, PEFT_EPI_RESID_F32, 0}, {300, 520, 264, PEFT_EPI_STORE, 0},
      {50432, 2304, 800, PEFT_EPI_STORE, 0}, {50432, 3072, 768, PEFT_EPI_GELU, 0}, {50432, 768, 3072, PEFT_EPI_RESID_F32, 0},
      {50432, 768, 768, PEFT_EPI_STORE, 0}, {50432, 768, 2336, PEFT_EPI_STORE, 0}, {50432, 3072, 768, PEFT_EPI_DGELU, 0},
      {50432, 768, 768, PEFT_EPI_RESID_F32, 0}, {50432, 768, 3072, PEFT_EPI_STORE, 0}, {32896, 4096, 1024, PEFT_EPI_QUICKGELU, 0},
      {8192, 8192, 8192, PEFT_EPI_STORE, 0},
  };
  int fails = 0;
  for (auto& c : cases) fails += run_case(c, timeit);
  printf("%s: %d failing cases\n", fails ? "FAILED" : "ALL PASS", fails);
  return fails ? 1 : 0;
}
