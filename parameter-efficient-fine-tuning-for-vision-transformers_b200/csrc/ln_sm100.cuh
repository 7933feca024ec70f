// LayerNorm forward / backward, one warp per token row, 128-bit vectorised HBM access.
//   forward : x (fp32 residual stream, pitch ldx) -> y (bf16 GEMM operand or fp32 stream, pitch ldy), mean, rstd
//   backward: g_out(bf16) = g_in(bf16) + LN'(dy)   — the residual-gradient add is fused; optional dgamma/dbeta
//             (only the Houlsby/Compacter adapter_norm_before is trainable; backbone gammas are frozen -> skipped).
// HBM-bound: the roofline is bytes moved / 6.55 TB/s.
#pragma once
#include "ptx_sm100.cuh"

namespace peft {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <typename T>
__device__ __forceinline__ void load4(const T* p, float (&v)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float (&v)[4]) {
  const float4 f = *reinterpret_cast<const float4*>(p);
  v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[4]) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = ptx::unpack_bf16x2(u.x), b = ptx::unpack_bf16x2(u.y);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
template <typename T>
__device__ __forceinline__ void store4(T* p, const float (&v)[4]);
template <>
__device__ __forceinline__ void store4<float>(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[4]) {
  uint2 u;
  u.x = ptx::pack_bf16x2(v[0], v[1]);
  u.y = ptx::pack_bf16x2(v[2], v[3]);
  *reinterpret_cast<uint2*>(p) = u;
}

// NV = D / 128: each lane owns NV groups of 4 consecutive columns (column = (i*32 + lane)*4).
template <int NV, typename InT, typename OutT>
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const InT* __restrict__ x, long long ldx, const float* __restrict__ gamma, const float* __restrict__ beta,
              OutT* __restrict__ y, long long ldy, float* __restrict__ mean_out, float* __restrict__ rstd_out, int M,
              float eps, const float* __restrict__ add_in = nullptr, long long ldai = 0, float* __restrict__ add_out = nullptr,
              long long ldao = 0) {
  pdl_trigger();                     // PDL (common.cuh): CTAs become resident during the predecessor's tail ...
  pdl_wait();                        // ... and read nothing before it has completed
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int D = NV * 128;
  float g[NV][4], b[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    load4<float>(gamma + (i * 32 + lane) * 4, g[i]);
    load4<float>(beta + (i * 32 + lane) * 4, b[i]);
  }
  // two rows per warp iteration, all loads issued before the first reduction: twice the bytes in flight per warp (the
  // kernel is latency-bound, not issue-bound)
  const long long stride = (long long)gridDim.x * warps_per_block;
  for (long long row0 = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); row0 < M; row0 += 2 * stride) {
    const long long row1 = row0 + stride;
    const bool two = row1 < M;
    float v[2][NV][4];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      load4<InT>(x + row0 * ldx + (i * 32 + lane) * 4, v[0][i]);
      if (two) load4<InT>(x + row1 * ldx + (i * 32 + lane) * 4, v[1][i]);
    }
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      if (rr == 1 && !two) break;
      const long long row = rr == 0 ? row0 : row1;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        s += v[rr][i][0] + v[rr][i][1] + v[rr][i][2] + v[rr][i][3];
        if (add_out != nullptr) {   // side output: add_out = x + add_in (the residual sum the adapter's up-projection lands on)
          float a[4];
          load4<float>(add_in + row * ldai + (i * 32 + lane) * 4, a);
#pragma unroll
          for (int j = 0; j < 4; ++j) a[j] += v[rr][i][j];
          store4<float>(add_out + row * ldao + (i * 32 + lane) * 4, a);
        }
      }
      const float mean = warp_sum(s) * (1.0f / D);
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float d = v[rr][i][j] - mean;
          q += d * d;
        }
      const float rstd = rsqrtf(warp_sum(q) * (1.0f / D) + eps);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = (v[rr][i][j] - mean) * rstd * g[i][j] + b[i][j];
        store4<OutT>(y + row * ldy + (i * 32 + lane) * 4, o);
      }
      if (lane == 0) {
        if (mean_out) mean_out[row] = mean;
        if (rstd_out) rstd_out[row] = rstd;
      }
    }
  }
}

// dx = rstd * (gamma*dy - mean(gamma*dy) - xhat * mean(gamma*dy*xhat)),  xhat = (x - mean) * rstd
// g_out = (g_in ? g_in : 0) + dx.   With kParamGrad: dgamma += sum_rows dy*xhat, dbeta += sum_rows dy (fp32 atomics,
// one per column per CTA after an in-CTA shared-memory reduction; the host launches this variant with few CTAs).
template <int NV, typename DyT, typename XT, typename GT, bool kParamGrad>
__global__ void __launch_bounds__(256)
ln_bwd_kernel(const DyT* __restrict__ dy, long long lddy, const XT* __restrict__ x, long long ldx,
              const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const float* __restrict__ gamma,
              const GT* __restrict__ g_in, long long ldgi, GT* __restrict__ g_out, long long ldgo,
              float* __restrict__ dgamma, float* __restrict__ dbeta, int M,
              __nv_bfloat16* __restrict__ g_copy = nullptr, long long ldgc = 0) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps_per_block = blockDim.x >> 5;
  const int D = NV * 128;
  float g[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i) load4<float>(gamma + (i * 32 + lane) * 4, g[i]);
  float ag[NV][4], ab[NV][4];
  if constexpr (kParamGrad) {
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) ag[i][j] = ab[i][j] = 0.f;
  }
  for (long long row = (long long)blockIdx.x * warps_per_block + warp; row < M;
       row += (long long)gridDim.x * warps_per_block) {
    const float mean = mean_in[row], rstd = rstd_in[row];
    float d[NV][4], xh[NV][4], o[NV][4];
    float s1 = 0.f, s2 = 0.f;
    // every load of the row is issued up front — the residual-gradient input too, although it is only needed after the two
    // warp reductions: one HBM round trip per row instead of two (the kernel is latency-bound at 16 warps per SM)
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      load4<DyT>(dy + row * lddy + (i * 32 + lane) * 4, d[i]);
      load4<XT>(x + row * ldx + (i * 32 + lane) * 4, xh[i]);
      if (g_in != nullptr) load4<GT>(g_in + row * ldgi + (i * 32 + lane) * 4, o[i]);
      else o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        xh[i][j] = (xh[i][j] - mean) * rstd;
        if constexpr (kParamGrad) {
          ag[i][j] += d[i][j] * xh[i][j];
          ab[i][j] += d[i][j];
        }
        d[i][j] *= g[i][j];
        s1 += d[i][j];
        s2 += d[i][j] * xh[i][j];
      }
    }
    s1 = warp_sum(s1) * (1.0f / D);
    s2 = warp_sum(s2) * (1.0f / D);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
#pragma unroll
      for (int j = 0; j < 4; ++j) o[i][j] += rstd * (d[i][j] - s1 - xh[i][j] * s2);
      store4<GT>(g_out + row * ldgo + (i * 32 + lane) * 4, o[i]);
      if (g_copy != nullptr) store4<__nv_bfloat16>(g_copy + row * ldgc + (i * 32 + lane) * 4, o[i]);
    }
  }
  if constexpr (kParamGrad) {
    __shared__ float red[2][NV * 128];
    for (int c = threadIdx.x; c < 2 * D; c += blockDim.x) (&red[0][0])[c] = 0.f;
    __syncthreads();
    // warp by warp, each lane adding into its own columns: no shared-memory float atomics (they compile to CAS loops)
    for (int wsel = 0; wsel < warps_per_block; ++wsel) {
      if (warp == wsel) {
#pragma unroll
        for (int i = 0; i < NV; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            red[0][(i * 32 + lane) * 4 + j] += ag[i][j];
            red[1][(i * 32 + lane) * 4 + j] += ab[i][j];
          }
      }
      __syncthreads();
    }
    for (int c = threadIdx.x; c < D; c += blockDim.x) {
      atomicAdd(dgamma + c, red[0][c]);
      atomicAdd(dbeta + c, red[1][c]);
    }
  }
}

// ------------------------------------------------------------------------------------------------ patch embedding
// images fp32 [B,3,H,W] -> bf16 patch rows [B*P, 3*p*p], column = c*p*p + i*p + j (matches conv weight flatten).
// replaces the unfold half of nn.Conv2d(stride=kernel=p): FS/model.py:1203, FS/kadaptation.py:300-307.
template <int VEC>
__global__ void im2col_patch_kernel(const float* __restrict__ img, __nv_bfloat16* __restrict__ out, long long ldo, int B,
                                    int HW, int p) {
  const int gp = HW / p;                // patches per side
  const int K = 3 * p * p;
  const long long total = (long long)B * gp * gp * K / VEC;  // VEC consecutive j per thread (p % VEC == 0)
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long e = t * VEC;
    const int k = (int)(e % K);
    const long long row = e / K;
    const int c = k / (p * p), i = (k / p) % p, j = k % p;
    const int px = (int)(row % gp), py = (int)((row / gp) % gp);
    const int b = (int)(row / (gp * gp));
    const float* src = img + (((long long)b * 3 + c) * HW + (py * p + i)) * HW + px * p + j;
    __nv_bfloat16* dst = out + row * ldo + k;
    if (VEC == 4) {
      const float4 f = *reinterpret_cast<const float4*>(src);
      uint2 u;
      u.x = ptx::pack_bf16x2(f.x, f.y);
      u.y = ptx::pack_bf16x2(f.z, f.w);
      *reinterpret_cast<uint2*>(dst) = u;
    } else {
      const float2 f = *reinterpret_cast<const float2*>(src);
      *reinterpret_cast<uint32_t*>(dst) = ptx::pack_bf16x2(f.x, f.y);
    }
  }
}

// tokens fp32 [B, L, D]: row 0 = cls + pos[0]; row 1+t = patch_embed[b*P+t] (bf16, conv bias already added) + pos[1+t]
// replaces torch.cat(cls, x) + pos_embed: FS/kadaptation.py:524-529, FS/model.py:1206-1207.
__global__ void assemble_tokens_kernel(const __nv_bfloat16* __restrict__ pe, const float* __restrict__ cls,
                                       const float* __restrict__ pos, float* __restrict__ x, int B, int L, int D) {
  const long long total = (long long)B * L * D / 4;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long e = t * 4;
    const int d = (int)(e % D);
    const long long row = e / D;
    const int l = (int)(row % L);
    const long long b = row / L;
    float v[4], pz[4];
    load4<float>(pos + (long long)l * D + d, pz);
    if (l == 0) load4<float>(cls + d, v);
    else load4<__nv_bfloat16>(pe + (b * (L - 1) + (l - 1)) * D + d, v);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] += pz[j];
    store4<float>(x + e, v);
  }
}

}  // namespace peft
