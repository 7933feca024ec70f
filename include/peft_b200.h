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

#ifdef __cplusplus
}
#endif
#endif /* PEFT_B200_H */
