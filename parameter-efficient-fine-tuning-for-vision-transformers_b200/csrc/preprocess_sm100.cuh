// CLIP input pipeline on the GPU: Resize(bicubic, antialias) -> CenterCrop -> ToTensor -> Normalize of
// few_shot/vision_benchmark/evaluation/clip_load.py:80-87, in the exact 8-bit fixed-point arithmetic of Pillow's two-pass
// resampler (the reference applies the transform to PIL images): per output pixel  clip8((2^21 + sum_x pixel[x] * k[x]) >> 22)
// with 22-bit coefficients, horizontal pass first (rounded to uint8), then the vertical pass.  Only the rows / columns the
// centre crop needs are computed.  HBM-bound byte work: coalesced over the interleaved RGB bytes of a row.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace peft {

__device__ __forceinline__ uint8_t pp_clip8(int v) {
  v >>= 22;
  return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// tmp[r - r0][ox][c] for r in [r0, r1), ox in [0, n), c in [0, 3): horizontal resampling of input row r at output column left + ox
__global__ void __launch_bounds__(256) pp_horizontal_kernel(const uint8_t* __restrict__ img, int W, const int* __restrict__ hb,
                                                            const int* __restrict__ hk, int ksize, int left, int n, int r0, int r1,
                                                            uint8_t* __restrict__ tmp) {
  const long long total = (long long)(r1 - r0) * n * 3;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(t % 3);
    const int ox = (int)((t / 3) % n);
    const int r = r0 + (int)(t / (3LL * n));
    const int xo = left + ox;
    const int xmin = hb[2 * xo], cnt = hb[2 * xo + 1];
    const int* k = hk + (long long)xo * ksize;
    const uint8_t* src = img + ((long long)r * W + xmin) * 3 + c;
    int ss = 1 << 21;
    for (int x = 0; x < cnt; ++x) ss += (int)src[3 * x] * k[x];
    tmp[t] = pp_clip8(ss);
  }
}

// out[c][oy][ox] = (clip8(vertical resampling of tmp) / 255 - mean[c]) / std[c]      (ToTensor, then Normalize: same op order)
__global__ void __launch_bounds__(256) pp_vertical_kernel(const uint8_t* __restrict__ tmp, const int* __restrict__ vb,
                                                          const int* __restrict__ vk, int ksize, int top, int n, int r0,
                                                          const float* __restrict__ mean, const float* __restrict__ stdv,
                                                          float* __restrict__ out) {
  const long long total = (long long)n * n * 3;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(t % 3);
    const int ox = (int)((t / 3) % n);
    const int oy = (int)(t / (3LL * n));
    const int yo = top + oy;
    const int ymin = vb[2 * yo], cnt = vb[2 * yo + 1];
    const int* k = vk + (long long)yo * ksize;
    const uint8_t* src = tmp + ((long long)(ymin - r0) * n + ox) * 3 + c;
    int ss = 1 << 21;
    for (int y = 0; y < cnt; ++y) ss += (int)src[(long long)y * n * 3] * k[y];
    const float v = (float)pp_clip8(ss) / 255.0f;
    out[((long long)c * n + oy) * n + ox] = (v - mean[c]) / stdv[c];
  }
}

}  // namespace peft
