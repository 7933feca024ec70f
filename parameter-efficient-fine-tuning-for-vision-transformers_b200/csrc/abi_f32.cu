// C-ABI of the fp32 verification mode: plain CUDA-core counterparts of the bf16 / tensor-core entry points.
#include "common.cuh"
#include "f32_sm100.cuh"

using namespace peft;

namespace {
inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
inline int grid_for(long long total, int threads) {
  long long blocks = (total + threads - 1) / threads;
  const long long cap = (long long)sm_count() * 16;
  return (int)(blocks > cap ? cap : (blocks < 1 ? 1 : blocks));
}
}  // namespace

extern "C" int peft_gemm_f32(const float* a, long long lda, const float* b, long long ldb, int M, int N, int K,
                             const float* bias, float* out, long long ldo, float* out2, long long ldo2, const float* aux,
                             long long ldaux, int epilogue, void* stream) {
  return peft_gemm_f32_rs(a, lda, b, ldb, M, N, K, bias, out, ldo, out2, ldo2, aux, ldaux, epilogue, nullptr, stream);
}

extern "C" int peft_gemm_f32_rs(const float* a, long long lda, const float* b, long long ldb, int M, int N, int K,
                                const float* bias, float* out, long long ldo, float* out2, long long ldo2, const float* aux,
                                long long ldaux, int epilogue, const float* rowscale, void* stream) {
  PEFT_REQUIRE(M > 0 && N > 0 && K > 0 && a && b && out, "peft_gemm_f32: bad shape M=%d N=%d K=%d or null operand", M, N, K);
  PEFT_REQUIRE(epilogue >= 0 && epilogue <= 11, "peft_gemm_f32: unknown epilogue %d", epilogue);
  const bool dual = epilogue == 1 || epilogue == 2 || epilogue == 9;
  const bool need_aux = epilogue == 3 || epilogue == 4 || epilogue == 5 || epilogue == 8 || epilogue == 10 || epilogue == 11;
  PEFT_REQUIRE(!dual || out2, "peft_gemm_f32: epilogue %d needs out2", epilogue);
  PEFT_REQUIRE(!need_aux || aux, "peft_gemm_f32: epilogue %d needs aux", epilogue);
  f32::GemmArgs p{};
  p.a = a; p.sa_m = lda; p.sa_k = 1; p.b = b; p.sb_n = ldb; p.sb_k = 1; p.M = M; p.N = N; p.K = K; p.bias = bias;
  p.out = out; p.ldo = ldo; p.out2 = out2; p.ldo2 = ldo2; p.aux = need_aux ? aux : nullptr; p.ldaux = ldaux; p.epi = epilogue;
  p.kchunk = K;
  p.rowscale = rowscale;
  f32::gemm_kernel<<<dim3((N + 63) / 64, (M + 63) / 64, 1), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(p);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_tn_reduce_f32(const float* x, long long ldx, int I, const float* y, long long ldy, int J, float* out,
                                  long long ldout, int M, void* stream) {
  PEFT_REQUIRE(I > 0 && J > 0 && M > 0 && x && y && out, "peft_tn_reduce_f32: bad shape I=%d J=%d M=%d", I, J, M);
  f32::GemmArgs p{};
  p.a = x; p.sa_m = 1; p.sa_k = ldx; p.b = y; p.sb_n = 1; p.sb_k = ldy; p.M = I; p.N = J; p.K = M;
  p.out = out; p.ldo = ldout; p.epi = f32::kEpiAtomicAdd;
  p.kchunk = 1024;
  f32::gemm_kernel<<<dim3((J + 63) / 64, (I + 63) / 64, (M + p.kchunk - 1) / p.kchunk), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(p);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_attention_fwd_f32(const float* qkv, long long ldqkv, float* o, long long ldo, float* lse, int B, int H,
                                      int L, float scale, void* stream) {
  return peft_attention_fwd_f32_ex(qkv, ldqkv, o, ldo, lse, B, H, L, scale, 0, stream);
}

extern "C" int peft_attention_fwd_f32_ex(const float* qkv, long long ldqkv, float* o, long long ldo, float* lse, int B, int H,
                                         int L, float scale, int causal, void* stream) {
  PEFT_REQUIRE(B > 0 && H > 0 && L > 0 && qkv && o && lse, "peft_attention_fwd_f32: bad shape B=%d H=%d L=%d", B, H, L);
  PEFT_REQUIRE(al16(qkv) && al16(o) && ldqkv % 4 == 0 && ldo % 4 == 0 && ldqkv >= 3ll * H * 64, "peft_attention_fwd_f32: 16-byte aligned rows required");
  f32::attn_fwd_kernel<<<dim3((L + 127) / 128, B * H), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(qkv, ldqkv, o, ldo, lse, B, H, L, scale, causal);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_attention_bwd_f32(const float* qkv, long long ldqkv, const float* o, long long ldo, const float* d_o,
                                      long long lddo, const float* lse, float* delta, float* dqkv, long long lddqkv, int B,
                                      int H, int L, float scale, void* stream) {
  PEFT_REQUIRE(B > 0 && H > 0 && L > 0 && qkv && o && d_o && lse && delta && dqkv, "peft_attention_bwd_f32: bad shape or null operand");
  PEFT_REQUIRE(al16(qkv) && al16(o) && al16(d_o) && al16(dqkv) && ldqkv % 4 == 0 && ldo % 4 == 0 && lddo % 4 == 0 && lddqkv % 4 == 0,
               "peft_attention_bwd_f32: 16-byte aligned rows required");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const dim3 grid((L + 127) / 128, B * H);
  f32::attn_bwd_dq_kernel<<<grid, 128, 0, s>>>(qkv, ldqkv, o, ldo, d_o, lddo, lse, delta, dqkv, lddqkv, B, H, L, scale);
  PEFT_LAUNCH_CHECK();
  f32::attn_bwd_dkv_kernel<0><<<grid, 128, 0, s>>>(qkv, ldqkv, d_o, lddo, lse, delta, dqkv, lddqkv, B, H, L, scale);
  PEFT_LAUNCH_CHECK();
  f32::attn_bwd_dkv_kernel<1><<<grid, 128, 0, s>>>(qkv, ldqkv, d_o, lddo, lse, delta, dqkv, lddqkv, B, H, L, scale);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_clip_delta_f32(float* qkv, long long ldqkv, float* delta, long long ldd, int B, int H, int L,
                                   float mult_q, float mult_v, int backward, void* stream) {
  PEFT_REQUIRE(B > 0 && H > 0 && L > 0 && qkv && delta && al16(qkv) && al16(delta) && ldqkv % 4 == 0 && ldd % 4 == 0,
               "peft_clip_delta_f32: bad shape or unaligned operand");
  const int blocks = grid_for((long long)B * H * L * 32, 256);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (backward) f32::clip_delta_kernel<true><<<blocks, 256, 0, s>>>(qkv, ldqkv, delta, ldd, B, H, L, mult_q, mult_v);
  else f32::clip_delta_kernel<false><<<blocks, 256, 0, s>>>(qkv, ldqkv, delta, ldd, B, H, L, mult_q, mult_v);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_im2col_patches_f32(const float* images, float* patches, long long ldp, int B, int img, int patch, void* stream) {
  PEFT_REQUIRE(B > 0 && img > 0 && patch > 0 && img % patch == 0 && images && patches && ldp >= 3ll * patch * patch,
               "peft_im2col_patches_f32: bad shape B=%d img=%d patch=%d", B, img, patch);
  const long long total = (long long)B * (img / patch) * (img / patch) * 3 * patch * patch;
  f32::im2col_kernel<<<grid_for(total, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(images, patches, ldp, B, img, patch);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_assemble_tokens_f32(const float* patch_embed, const float* cls, const float* pos, float* x, int B, int L,
                                        int D, void* stream) {
  PEFT_REQUIRE(B > 0 && L > 1 && D > 0 && patch_embed && cls && pos && x, "peft_assemble_tokens_f32: bad shape or null operand");
  f32::assemble_kernel<<<grid_for((long long)B * L * D, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(patch_embed, cls, pos, x, B, L, D);
  PEFT_LAUNCH_CHECK();
  return 0;
}

extern "C" int peft_add_f32(const float* a, const float* b, float* out, long long n, void* stream) {
  PEFT_REQUIRE(n > 0 && a && b && out, "peft_add_f32: bad size or null operand");
  f32::add_kernel<<<grid_for(n, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, b, out, n);
  PEFT_LAUNCH_CHECK();
  return 0;
}
