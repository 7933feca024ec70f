/* libpeftb200.so — C ABI of the B200-native PEFT-ViT hot path.
 *
 * The reference (jkooy/Parameter-efficient-Fine-tuning-for-Vision-Transformers) is pure Python/PyTorch and has no
 * FFI; its seam is the nn.Module forward/backward contract.  Each entry point below replaces the
 * ATen calls made by one reference function; the citation after "replaces:" is relative to the reference root,
 * FS = few_shot/vision_benchmark/evaluation, FL = full_shot/main.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller (the torch allocator); the library never frees them;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); no implicit synchronisation;
 *   - return value 0 = success, negative = peft_status; peft_last_error() gives the thread-local message;
 *   - bf16 tensors are row-major with an explicit row pitch ("ld", in elements) so callers can hand in column
 *     slices of wider buffers (the [x | p] K-extended operand of the adapted linear);
 *   - no torch types, no C++ types, no exceptions cross this boundary.
 */
#ifndef PEFT_B200_H
#define PEFT_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  PEFT_OK = 0,
  PEFT_ERR_INVALID = -1,   /* bad shape / alignment / enum */
  PEFT_ERR_CUDA = -2,      /* a CUDA runtime or driver call failed */
  PEFT_ERR_UNSUPPORTED = -3
} peft_status;

/* epilogue selector of peft_gemm_bf16 (values mirror csrc/gemm_sm100.cuh) */
typedef enum {
  PEFT_EPI_STORE = 0, PEFT_EPI_GELU = 1, PEFT_EPI_QUICKGELU = 2, PEFT_EPI_RESID_F32 = 3, PEFT_EPI_DGELU = 4,
  PEFT_EPI_DQUICKGELU = 5, PEFT_EPI_STORE_F32 = 6, PEFT_EPI_RELU = 7, PEFT_EPI_DRELU = 8, PEFT_EPI_GELU_TANH = 9,
  PEFT_EPI_DGELU_TANH = 10, PEFT_EPI_RESID_BF16 = 11
} peft_epilogue;

int peft_version(void);
/* copies the calling thread's last error message (NUL-terminated, truncated to n) */
void peft_last_error(char* buf, size_t n);
/* number of kernels this library has launched since load (all threads) — bench.py reports it as gpu_launches */
unsigned long long peft_launch_count(void);
/* launches per kernel family since load — 0: 1-CTA tcgen05 GEMM, 1: 2-CTA (cta_group::2) GEMM, 2: tcgen05 attention forward,
 * 3: tcgen05 attention backward, 4: legacy mma.sync attention, 5: token-major (TN) tcgen05 GEMM, 6: 4-CTA multicast GEMM.
 * Test hook: the parity suite asserts that the benchmarked shapes really ran on the kernels the bench times. */
unsigned long long peft_kernel_count(int kind);
/* tuning / test knob: which tensor-core GEMM kernel peft_gemm_bf16 uses — 0 = automatic (2-CTA cta_group::2 kernel when
 * the problem fills the 74 SM pairs, else the 1-CTA kernel), 1 = always 1-CTA, 2 = always 2-CTA. */
int peft_set_gemm_algo(int algo);
/* programmatic dependent launch (cudaLaunchAttributeProgrammaticStreamSerialization) of the hot-path kernels: the prologue of a
 * kernel (barrier init, TMEM allocation, tensor-map fetch) overlaps the tail of its predecessor on the stream; every such kernel
 * executes griddepcontrol.wait before its first global-memory access.  1 = on (default; PEFT_PDL=0 in the environment turns it
 * off), 0 = fully serialised launches (A/B measurements). */
int peft_set_pdl(int enabled);
/* stream-K scheduling of the 2-CTA GEMM (long contractions whose tile count leaves the last wave mostly empty, e.g. 75 tiles
 * on 74 SM pairs at M = 6304): 0 = off, 1 = automatic (default), 2 = whenever the shape allows it (tests).  Partial tiles
 * travel through ONE per-device workspace, so stream-K products of one device must be issued on one stream at a time. */
int peft_set_gemm_streamk(int mode);
/* The in-kernel diagnostics behind the next two calls (and the PEFT_GEMM_DBG / PEFT_ATTN3_FLAGS modes) exist only in a library built
 * with `make DIAG=1` (-DPEFT_GEMM_DIAG=1); the production build compiles them out — they cost 14 % of the GEMM kernel's code and
 * measurably lengthen short launches — and the two calls then fail with a message. */
/* diagnostics (tests/bench_ops.py gemm_stats): with PEFT_GEMM_STATS=1 in the environment the 2-CTA GEMM records, per CTA,
 * the cycles its TMA producer, MMA issuer and first epilogue warp spent waiting on their mbarriers during the LAST launch:
 * host_out[cta*8 + {0: producer wait-empty, 1: producer total, 2: issuer wait-full, 3: issuer wait-tmem-empty, 4: issuer
 * total, 5: epilogue wait-accumulator, 6: epilogue total}].  Synchronises the device. */
int peft_debug_gemm_stats(unsigned long long* host_out, int n_ctas);
/* same for the tcgen05 attention forward (PEFT_ATTN_STATS=1): host_out[(cta*2 + warpgroup)*8 + {0: wait S, 1: max pass,
 * 2: exp pass, 3: wait O, 4: O epilogue + store, 5: total}] of the first softmax warp of each warpgroup. */
int peft_debug_attn_stats(unsigned long long* host_out, int n_ctas);
/* diagnostics: one CTA issues `reps` tcgen05.mma (M = 128, N = n, K = 16, bf16; ts != 0: A operand from TMEM) that rotate over
 * `nacc` accumulators (nacc = 1: a fully dependent chain); host_out[0] = cycles to issue them, host_out[1] = cycles until the
 * last one has completed.  Measures the dependent-accumulate latency that paces the attention kernels. */
int peft_debug_mma_probe(int n, int nacc, int ts, int reps, unsigned long long* host_out);
/* same for the attention kernels: 0 = automatic (tcgen05/TMEM kernels for L <= 272; backward: the software-pipelined pair of
 * attention3_bwd_sm100.cuh for L <= 64, the attention2 pair above), 1 = mma.sync kernels, 2 = tcgen05 with the attention2 backward
 * for every L, 3 = tcgen05 with the pipelined backward for every L <= 272.  Env PEFT_ATTN_ALGO sets the initial value. */
int peft_set_attention_algo(int algo);

/* C[M,N] = A[M,K] * B[N,K]^T with fused epilogue, bf16 operands, fp32 accumulate in TMEM (tcgen05 + TMA).
 * replaces: F.linear / nn.Linear on the hot path —
 *   qkv          FS/kadaptation.py:133,202   FS/model.py:382 (in_proj_weight/in_proj_bias)
 *   out-proj     FS/kadaptation.py:236       FS/model.py:1003
 *   mlp fc1/fc2  FS/kadaptation.py:72-78     FS/model.py:1143-1147 (c_fc, QuickGELU, c_proj)
 *   + the adapter term of FS/kadaptation.py:211-219 (matvec_product) / FS/lora.py adapter_forward as extra K columns
 *   + their autograd backward (dX only; frozen weights get no dW).
 * a: [M,K] pitch lda; b: [N,K] pitch ldb; bias fp32 [N] or NULL; out/out2/aux per peft_epilogue. */
int peft_gemm_bf16(const void* a, long long lda, const void* b, long long ldb, int M, int N, int K,
                   const float* bias, void* out, long long ldo, void* out2, long long ldo2,
                   const void* aux, long long ldaux, int epilogue, void* stream);
/* same, with (acc + bias) multiplied by rowscale[row] (fp32 [M], NULL = none) before the rest of the epilogue:
 * the per-sample stochastic-depth factor keep/keep_prob of `x + self.drop_path(branch)` (timm DropPath; FS/kadaptation.py:263,
 * 276-277, full_shot yaml DROP_PATH_RATE 0.1) folded into the residual epilogue (forward) and into the first dX product
 * of the branch (backward). */
int peft_gemm_bf16_rs(const void* a, long long lda, const void* b, long long ldb, int M, int N, int K,
                      const float* bias, void* out, long long ldo, void* out2, long long ldo2,
                      const void* aux, long long ldaux, int epilogue, const float* rowscale, void* stream);
/* Allocation-free variant of peft_gemm_bf16 (2-CTA tcgen05 kernel, stream-K whenever the shape allows it).  It has no
 * counterpart in the reference and was added: for callers that may not call cudaMalloc
 * (code under CUDA-graph capture without a warm-up call, tight serving loops, harnesses that forbid allocation).  The stream-K
 * fix-up scratch (one 256 x 256 fp32 partial-tile slot per SM pair + per-warp flags) lives in a file-scope
 * `static __device__ __align__(256)` array of 80 slots = 20 MiB (+ 5 KiB of flags) that every GPU loading the library holds
 * until exit; the call performs no cudaMalloc / cudaFree / allocating library call.  Limits: devices with more than 80 SM pairs
 * (160 SMs) and products of more than 2^24 output tiles are rejected with PEFT_ERR_INVALID and a message.  All calls on a device
 * share the one array (it keeps the last call's partial sums; the flags are cleared on the stream before every launch): calls
 * whose kernels overlap on the GPU race — issue them on one stream.  Same arguments, epilogues and results as peft_gemm_bf16. */
int peft_gemm_bf16_noalloc(const void* a, long long lda, const void* b, long long ldb, int M, int N, int K,
                           const float* bias, void* out, long long ldo, void* out2, long long ldo2,
                           const void* aux, long long ldaux, int epilogue, void* stream);

/* LayerNorm over the last dimension, one warp per token row, D % 128 == 0.
 * replaces: nn.LayerNorm / CLIP LayerNorm forward — FS/kadaptation.py:274-276,538 (eps 1e-6), FS/model.py:156-162,1156-1157,1208,1214 (eps 1e-5).
 * x: fp32 (x_is_bf16=0) or bf16 rows of pitch ldx; y: bf16 GEMM operand (y_is_f32=0) or fp32; mean/rstd: fp32 [M] or NULL. */
int peft_layernorm_fwd(const void* x, long long ldx, int x_is_bf16, const float* gamma, const float* beta,
                       void* y, long long ldy, int y_is_f32, float* mean, float* rstd, int M, int D, float eps,
                       void* stream);
/* y(bf16) = LN(x fp32) and, in the same pass, add_out(fp32) = x + add_in(fp32).
 * replaces: Adapter / HyperComplexAdapter's adapter_norm_before plus the residual adds of
 * `x + adapter(mlp(ln_2(x)))` — FS/adapter_model.py:230-232,330-334, FS/compacter_model.py:393-425,477-479. */
int peft_layernorm_fwd_add(const float* x, long long ldx, const float* gamma, const float* beta, void* y, long long ldy,
                           float* mean, float* rstd, const float* add_in, long long ldai, float* add_out, long long ldao,
                           int M, int D, float eps, void* stream);
/* g_out(bf16) = (g_in ? g_in : 0) + dLN/dx(dy); optional dgamma/dbeta accumulation (fp32 atomics, caller zeroes them).
 * replaces: the autograd backward of the LayerNorms above, fused with the residual-gradient add of Block.forward
 * (FS/kadaptation.py:273-277, FS/model.py:1155-1158). */
int peft_layernorm_bwd(const void* dy, long long lddy, int dy_is_f32, const void* x, long long ldx, int x_is_bf16,
                       const float* mean, const float* rstd, const float* gamma, const void* g_in, long long ldgi,
                       void* g_out, long long ldgo, float* dgamma, float* dbeta, int M, int D, void* stream);

/* same with an fp32 residual-gradient stream: g_out(fp32) = (g_in ? g_in : 0) + dLN/dx(dy) (g_in may alias g_out);
 * g_copy (bf16, optional) receives a rounded copy of g_out — the A operand of the next dX product. */
int peft_layernorm_bwd_f32g(const void* dy, long long lddy, int dy_is_f32, const void* x, long long ldx, int x_is_bf16,
                            const float* mean, const float* rstd, const float* gamma, const float* g_in, long long ldgi,
                            float* g_out, long long ldgo, void* g_copy, long long ldgc, float* dgamma, float* dbeta, int M,
                            int D, void* stream);

/* softmax(q k^T * scale) v for every (batch, head), head_dim 64, reading the packed projection buffer
 * qkv[B*L, >=3*H*64] (q | k | v column blocks, token row = b*L + l) and writing o[B*L, H*64]; lse2[B*H*L] receives
 * rowmax*scale*log2e + log2(rowsum) for the backward.
 * replaces: FS/kadaptation.py:202-205,223-235 and FS/model.py:907-918,990-1002 (head split, bmm, softmax, bmm, merge). */
int peft_attention_fwd(const void* qkv, long long ldqkv, void* o, long long ldo, float* lse2, int B, int H, int L,
                       float scale, void* stream);
/* dqkv[B*L, >=3*H*64] (dq | dk | dv) from d_o; delta[B*H*L] is scratch.  replaces: autograd of the above. */
int peft_attention_bwd(const void* qkv, long long ldqkv, const void* o, long long ldo, const void* d_o, long long lddo,
                       const float* lse2, float* delta, void* dqkv, long long lddqkv, int B, int H, int L, float scale,
                       void* stream);

/* images fp32 [B,3,img,img] -> bf16 patch rows [B*(img/patch)^2, 3*patch*patch] of pitch ldp >= 3*patch*patch, ldp % 8 == 0
 * (columns beyond 3*patch*patch are left untouched: the caller keeps them zero, e.g. 588 -> 592 for patch 14).  The
 * unfold half of the stride=kernel conv of FS/model.py:1203 / FS/kadaptation.py:300-307; the GEMM half is peft_gemm_bf16. */
int peft_im2col_patches(const float* images, void* patches, long long ldp, int B, int img, int patch, void* stream);
/* x fp32 [B,L,D]: row 0 = cls + pos[0], row 1+t = patch_embed[b*(L-1)+t] + pos[1+t]
 * replaces: FS/kadaptation.py:524-529, FS/model.py:1206-1207. */
int peft_assemble_tokens(const void* patch_embed, const float* cls, const float* pos, float* x, int B, int L, int D,
                         void* stream);

/* out[I, J] (fp32, pitch ldout) += sum_m x[m, i] * y[m, j]; I in {16,32,64}, J % 64 == 0.  The adapter-gradient
 * reductions over all tokens: replaces the autograd dW of the LoRA / Kronecker / Houlsby / Compacter factors
 * (FS/model.py:754-762, FS/kadaptation.py:81-115,211-219, FS/adapter_model.py:265-267). */
int peft_tn_reduce(const void* x, long long ldx, int I, const void* y, long long ldy, int J, float* out,
                   long long ldout, int M, void* stream);

/* CLIP-dialect placement of the q/v adapter update (SURVEY.md §8a Q1 + Q2): the reference reshapes the [L,B,C] update
 * RAW to [B*H, L, hd] before adding it to the per-head q / v (FS/model.py:981-986, FS/lora_model.py:730-735) and q was
 * divided by sqrt(hd) beforehand (FS/model.py:967).  With token rows ordered b*L + l:
 *   backward == 0:  qkv[(b,l), h] (q and v blocks) += mult * delta[(b_s,l_s), slot]       (in place on qkv)
 *   backward == 1:  delta[(b_s,l_s), slot] = mult * qkv[(b,l), h]                          (qkv holds dq|dk|dv)
 * where g = (b*H+h)*L + l, (r, slot) = divmod(g, H), (l_s, b_s) = divmod(r, B); delta is [B*L, 2*H*64] (q | v). */
int peft_clip_delta(void* qkv, long long ldqkv, void* delta, long long ldd, int B, int H, int L, float mult_q,
                    float mult_v, int backward, void* stream);

/* out[M, N] (fp32, contiguous) = x[T, M]^T y[T, N]: the weight-gradient product dW = dY^T X taken straight from the
 * token-major activations (MN-major tcgen05 operands, no transposes), contraction over the T tokens split across tiles
 * when the output is small.  x, y: bf16, 16-byte aligned, pitches % 8 == 0; N % 4 == 0.
 * replaces: the autograd dW of nn.Linear in full fine-tuning / Fastfood mode (FL/intrinsic/fastfood.py:61-77 makes every
 * wrapped weight a function of V, so autograd needs dL/dW for all of them; FL/tools/finetune_{bias,attention}.py select
 * subsets), and dUp^T of the dense large-n KAdaptation update. */
int peft_gemm_tn_bf16(const void* x, long long ldx, const void* y, long long ldy, int M, int N, int T, float* out,
                      long long ldo, void* stream);
/* colsum[c] += sum_r in[r, c] (bf16 in, fp32 out, caller zeroes it): the bias gradient of a linear layer (autograd of
 * `+ bias`, FS/kadaptation.py:72-78,133,236) is the column sum of its dY.  16-byte aligned operand, pitch % 8 == 0. */
int peft_colsum_bf16(const void* in, long long ldin, int rows, int cols, float* colsum, void* stream);

/* out[c, r] = in[r, c], bf16 [rows, cols] (pitch ldin) -> [cols, rows] (pitch ldout).  Used by the per-step weight
 * re-pack of the full mode (W^T operands of the dX products from the bf16 copy of W); 16-byte aligned operands with
 * pitches % 8 == 0 take the shared-memory-free fast path, anything else (even pitches, 4-byte alignment) a tiled one. */
int peft_transpose_bf16(const void* in, long long ldin, int rows, int cols, void* out, long long ldout, void* stream);
/* same, and colsum[c] += sum_r in[r, c] in the same pass (fast path only) */
int peft_transpose_colsum_bf16(const void* in, long long ldin, int rows, int cols, void* out, long long ldout, float* colsum,
                               void* stream);

/* Large-n KAdaptation fold (the committed phm_dim = n = D, K = P = 1).
 * replaces: kronecker_product_einsum_batched(phm_rule, W).sum(0) of matvec_product (FS/kadaptation.py:81-115) for BOTH
 * projections in one pass over phm_rule (n^3 fp32, read once; HBM-bound):
 *   backward == 0:  out[p, j] = sum_i s[p, i] * rule[i, j]      s [2, n] (q, v),  out = H [2, n*n]
 *   backward == 1:  out[p, i] = sum_j dH[p, j] * rule[i, j]     dH [2, n*n],      out = ds [2, n]      (rule is frozen) */
int peft_phm_fold(const float* rule, const float* s_or_dH, float* out, int n, int backward, void* stream);

/* Small-n KAdaptation: the reshape-trick operands of ALL layers in one launch per direction (csrc/factors_sm100.cuh).
 * replaces: the stack / einsum / eye / zeros / cat chain (and its autograd) that builds, from the reference-named parameters
 * phm_rule A [dep,n,n,n] (frozen), W_left{1,2} u^s [dep,n,K,r], W_right{1,2} v^s [dep,n,r,P] (FS/kadaptation.py:81-115,143-187),
 *   Dn  [dep, 2nrn, nK]:  Dn[l,(s,i,r,c),(a,k)] = A_i[a,c] u^s_i[k,r]      UpT [dep, 2nrn, 3nP]: UpT[l,(s,i,r,c),(slot_s,c,p)] = v^s_i[r,p]
 * (s = 0: q, column slot 0; s = 1: v, column slot 2; zero elsewhere).
 *   backward == 0: a..d = wl1, wr1, wl2, wr2 (read), out0 = Dn, out1 = UpT (written)
 *   backward != 0: out0 = dDn, out1 = dUpT (read), a..d = dW_left1, dW_right1, dW_left2, dW_right2 (written, not accumulated) */
int peft_kron_factors(const float* rule, float* a, float* b, float* c, float* d, float* out0, float* out1, int dep, int n, int K,
                      int r, int P, int backward, void* stream);
/* ... and the gradient of a trainable phm_rule: drule[l,i,a,c] = sum_{s,r,k} W_left{s}[l,i,k,r] dDn[l,(s,i,r,c),(a,k)] (written). */
int peft_kron_factors_drule(const float* dDn, const float* wl1, const float* wl2, float* drule, int dep, int n, int K, int r,
                            void* stream);

/* Fastfood intrinsic-dimension projection over a SET of wrapped parameter tensors in one call.
 * replaces: FastfoodWrap.forward / fastfood_torched / fast_walsh_hadamard_torched and their autograd backward —
 * FL/intrinsic/fastfood.py:61-77,81-121,151-187.
 * Per tensor p (descriptor below): DD = numel, LL = 2^ll >= DD, coef = 1 / (divisor * sqrt(DD / LL)) with
 * divisor = sqrt(LL * sum(G^2)) (fastfood.py:124-148).  Device arrays, concatenated over tensors:
 *   pi_lo [sum LL] uint16 = Pi[j] mod min(LL, wn);  gg [sum LL] fp32 = G;  bsign [T, wn] fp32 = B[:wn] (zero padded);
 *   theta0 / theta / dtheta fp32 indexed by off_dd;  scratch [sum LL] fp32;  wu [T, wn] fp32 scratch;  counters [T + 1] int
 *   (per-tensor completion counters + one work-item ticket: items are taken by atomic ticket, never by blockIdx, so an item
 *   only ever waits on items that running CTAs already hold — forward progress is independent of block dispatch order).
 * wn = power of two >= d, <= 4096.  `items` is a device array of (tensor, tile | kind << 30) int pairs built by the
 * caller — kind 1 = "high" tile (tensors with ll > 13: tiles 0 .. LL/8192 - 1), kind 0 = "low" block (blocks
 * 0 .. ceil(DD/8192) - 1).  Forward lists a tensor's high tiles BEFORE its low blocks, backward its low blocks BEFORE its
 * high tiles; items of different tensors may interleave freely (the second pass waits for the first through `counters`;
 * the kernels are persistent cooperative launches taking items round-robin, so an item only ever waits on lower-numbered
 * ones).  off_ll and off_dd should be multiples of 64 elements: the DRAM streams are staged with 16-byte cp.async and the
 * forward drops the intermediate's L2 lines by 128-byte line; other offsets fall back to direct loads.  theta0 / dtheta
 * must be readable up to the next multiple of 4 elements past DD.  The forward leaves `scratch` undefined. */
typedef struct {
  long long off_ll;  /* element offset of this tensor in pi_lo / gg / scratch */
  long long off_dd;  /* element offset in theta0 / theta / dtheta */
  int DD;
  int ll;
  float coef;
  int reserved;
} peft_ff_tensor;
/* theta[off_dd + j] = theta0[off_dd + j] + coef * [H G Pi H B pad(V)][j],  j < DD, for every tensor */
int peft_fastfood_fwd(const float* V, int d, int wn, const peft_ff_tensor* tensors, int T, const int* items, int n_items,
                      const uint16_t* pi_lo, const float* gg, const float* bsign, const float* theta0, float* theta,
                      float* wu, float* scratch, int* counters, void* stream);
/* dV[k] += sum_p coef_p * B_p[k] * [H Pi^T G H pad(dtheta_p)][k],  k < d  (caller zeroes dV; wu is zeroed inside) */
int peft_fastfood_bwd(const float* dtheta, int d, int wn, const peft_ff_tensor* tensors, int T, const int* items,
                      int n_items, const uint16_t* pi_lo, const float* gg, const float* bsign, float* wu, float* scratch,
                      int* counters, float* dV, void* stream);

/* One-shot all-reduce (mean) of the flat fp32 gradient bucket over NVLink peer memory, fused with the SGD-momentum update of
 * the flat parameter replica and the zeroing of the bucket for the next step.
 * replaces: dist.all_reduce + 1/N scaling + torch.optim.SGD.step() + zero_grad of the data-parallel PEFT step; optimiser
 * semantics of few_shot/vision_benchmark/optim/build.py:120-159 (SGD: momentum mu, dampening 0, weight decay, no Nesterov),
 * stepped at few_shot/vision_benchmark/evaluation/kadaptation_clip.py:308-314.
 * peer_grads / peer_pads: DEVICE arrays of `world` pointers to every rank's bucket / signal pad (torch symmetric memory:
 * buffer_ptrs_dev / signal_pad_ptrs_dev); pad slots [pad_base, pad_base + 2 * blocks * world) of the uint32 pad are used
 * (blocks <= 32) and are left zero.  lr is a device scalar (LR schedules need no graph re-capture).  world == 1: plain fused
 * SGD + zeroing, no peer access.  n %% 4 == 0.  Every rank ends with bit-identical parameters (rank-ordered sums). */
int peft_allreduce_sgd(const void* const* peer_grads, void* const* peer_pads, int rank, int world, float* grad, float* param,
                       float* momentum, long long n, const float* lr, float mu, float weight_decay, int zero_grad,
                       int pad_base, int pad_slots, void* stream);

/* CLIP input pipeline on a decoded uint8 RGB image [H, W, 3] in device memory: Resize(n, BICUBIC) -> CenterCrop(n) -> ToTensor ->
 * Normalize(mean, std), bit-identical to the reference's torchvision-on-PIL transform.
 * replaces: `_transform(n_px)` of few_shot/vision_benchmark/evaluation/clip_load.py:80-87 (Pillow's two-pass 8-bit resampler:
 * 22-bit fixed-point coefficients, horizontal pass rounded to uint8, then vertical).  hb / vb: int32 [out, 2] (first input index,
 * count) and hk / vk: int32 [out, ksize] fixed-point coefficients of the horizontal / vertical axis (built by the host,
 * peft_b200/preprocess.py); top / left: crop origin in the resized image; rows [r0, r1) of the input are the ones the cropped
 * output needs; tmp: uint8 [(r1 - r0), n, 3] scratch; out: fp32 [3, n, n]. */
int peft_preprocess_clip(const uint8_t* img, int H, int W, const int* hb, const int* hk, int hksize, const int* vb, const int* vk,
                         int vksize, int top, int left, int n, int r0, int r1, const float* mean, const float* stdv, uint8_t* tmp,
                         float* out, void* stream);

/* ---- fp32 verification mode ("within 1e-5 relative for fp32 mode", BASELINE.json north_star) -------------------------
 * Plain CUDA-core fp32 counterparts of the entry points above: same operand layouts (row pitches in elements), same
 * epilogue semantics with every bf16 tensor replaced by fp32, IEEE expf/erff/tanhf.  The Python engine selects them
 * through the dtype of its buffers (EngineCfg.precision = "fp32"); peft_layernorm_fwd / _bwd_f32g already take fp32 on
 * both sides.  They replace the same reference lines as their bf16 twins and are not the throughput path. */
int peft_gemm_f32(const float* a, long long lda, const float* b, long long ldb, int M, int N, int K, const float* bias,
                  float* out, long long ldo, float* out2, long long ldo2, const float* aux, long long ldaux, int epilogue,
                  void* stream);
int peft_gemm_f32_rs(const float* a, long long lda, const float* b, long long ldb, int M, int N, int K, const float* bias,
                     float* out, long long ldo, float* out2, long long ldo2, const float* aux, long long ldaux, int epilogue,
                     const float* rowscale, void* stream);
/* out[I, J] += x[M, I]^T y[M, J], any I, J (split over M, fp32 atomics) */
int peft_tn_reduce_f32(const float* x, long long ldx, int I, const float* y, long long ldy, int J, float* out,
                       long long ldout, int M, void* stream);
/* lse[B*H*L] = rowmax + ln(rowsum) of the scaled scores (private to the fp32 kernels) */
int peft_attention_fwd_f32(const float* qkv, long long ldqkv, float* o, long long ldo, float* lse, int B, int H, int L,
                           float scale, void* stream);
/* same with an optional causal mask (causal != 0: query l sees keys j <= l) — the text tower of CLIP.encode_text, which the
 * trainers run once per task to initialise the classifier head from prompt embeddings (FS/model.py:1303-1327,
 * FS/kadaptation_clip.py:128-138); forward only. */
int peft_attention_fwd_f32_ex(const float* qkv, long long ldqkv, float* o, long long ldo, float* lse, int B, int H, int L,
                              float scale, int causal, void* stream);
int peft_attention_bwd_f32(const float* qkv, long long ldqkv, const float* o, long long ldo, const float* d_o, long long lddo,
                           const float* lse, float* delta, float* dqkv, long long lddqkv, int B, int H, int L, float scale,
                           void* stream);
int peft_clip_delta_f32(float* qkv, long long ldqkv, float* delta, long long ldd, int B, int H, int L, float mult_q,
                        float mult_v, int backward, void* stream);
int peft_im2col_patches_f32(const float* images, float* patches, long long ldp, int B, int img, int patch, void* stream);
int peft_assemble_tokens_f32(const float* patch_embed, const float* cls, const float* pos, float* x, int B, int L, int D,
                             void* stream);
/* out[i] = a[i] + b[i] (the residual add the bf16 path fuses into peft_layernorm_fwd_add) */
int peft_add_f32(const float* a, const float* b, float* out, long long n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PEFT_B200_H */
