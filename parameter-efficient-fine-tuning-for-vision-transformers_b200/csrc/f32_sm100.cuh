// fp32 verification mode (BASELINE.json north_star: "within 1e-5 relative for fp32 mode").
//
// Every tensor-core / bf16 kernel of the hot path has a plain fp32 CUDA-core counterpart here with the SAME operand
// layout and epilogue semantics, so the engine runs unchanged with fp32 buffers (EngineCfg.precision = "fp32").  These
// kernels exist to pin the algebra of the engine (K-extension, reshape trick, CLIP Q1/Q2 placement, adapter-gradient
// reductions) against the fp32 oracle at 1e-5; they are simple on purpose (one accumulation order, IEEE expf/erff/tanhf,
// no tensor cores) and are not the throughput path.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace peft {
namespace f32 {

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float dgelu_erf(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
  return cdf + x * 0.3989422804014327f * expf(-0.5f * x * x);
}
__device__ __forceinline__ float quickgelu(float x) { return x / (1.0f + expf(-1.702f * x)); }
__device__ __forceinline__ float dquickgelu(float x) {
  const float s = 1.0f / (1.0f + expf(-1.702f * x));
  return s * (1.0f + 1.702f * x * (1.0f - s));
}
__device__ __forceinline__ float gelu_tanh(float x) {
  return 0.5f * x * (1.0f + tanhf(0.7978845608028654f * (x + 0.044715f * x * x * x)));
}
__device__ __forceinline__ float dgelu_tanh(float x) {
  const float t = tanhf(0.7978845608028654f * (x + 0.044715f * x * x * x));
  const float du = 0.7978845608028654f * (1.0f + 3.0f * 0.044715f * x * x);
  return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * du;
}

// ---- GEMM: out[m, n] = epilogue(sum_k A(m, k) B(n, k) (+ bias[n])) with general element strides ----------------------
// A(m, k) = a[m * sa_m + k * sa_k], B(n, k) = b[n * sb_n + k * sb_k]: row-major [M,K] x [N,K]^T is (lda, 1, ldb, 1); the
// token reduction out[i, j] += sum_t x[t, i] y[t, j] is (1, ldx, 1, ldy) with gridDim.z splits over t and atomic adds.
constexpr int kEpiAtomicAdd = 100;

struct GemmArgs {
  const float* a; long long sa_m, sa_k;
  const float* b; long long sb_n, sb_k;
  int M, N, K;
  const float* bias;
  float* out; long long ldo;
  float* out2; long long ldo2;
  const float* aux; long long ldaux;
  int epi;
  int kchunk;   // K range per blockIdx.z
  const float* rowscale = nullptr;   // optional [M]: (acc + bias) *= rowscale[m] (stochastic depth)
};

__device__ __forceinline__ void gemm_epilogue(const GemmArgs& p, int m, int n, float acc) {
  float v = acc + (p.bias ? p.bias[n] : 0.f);
  if (p.rowscale) v *= p.rowscale[m];
  float* o = p.out + (long long)m * p.ldo + n;
  const float ax = p.aux ? p.aux[(long long)m * p.ldaux + n] : 0.f;
  switch (p.epi) {
    case 1: *o = v; p.out2[(long long)m * p.ldo2 + n] = gelu_erf(v); break;
    case 2: *o = v; p.out2[(long long)m * p.ldo2 + n] = quickgelu(v); break;
    case 9: *o = v; p.out2[(long long)m * p.ldo2 + n] = gelu_tanh(v); break;
    case 3: case 11: *o = ax + v; break;
    case 4: *o = v * dgelu_erf(ax); break;
    case 5: *o = v * dquickgelu(ax); break;
    case 10: *o = v * dgelu_tanh(ax); break;
    case 7: *o = v > 0.f ? v : 0.f; break;
    case 8: *o = ax > 0.f ? v : 0.f; break;
    case kEpiAtomicAdd: atomicAdd(o, v); break;
    default: *o = v; break;   // STORE / STORE_F32
  }
}

__global__ void __launch_bounds__(256) gemm_kernel(GemmArgs p) {
  __shared__ float As[16][68];
  __shared__ float Bs[16][68];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int row0 = blockIdx.y * 64, col0 = blockIdx.x * 64;
  const int kbeg = blockIdx.z * p.kchunk;
  const int kend = min(p.K, kbeg + p.kchunk);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  for (int k0 = kbeg; k0 < kend; k0 += 16) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int gk = k0 + lk + e, gm = row0 + lr, gn = col0 + lr;
      As[lk + e][lr] = (gm < p.M && gk < kend) ? p.a[(long long)gm * p.sa_m + (long long)gk * p.sa_k] : 0.f;
      Bs[lk + e][lr] = (gn < p.N && gk < kend) ? p.b[(long long)gn * p.sb_n + (long long)gk * p.sb_k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = row0 + ty * 4 + i, n = col0 + tx * 4 + j;
      if (m < p.M && n < p.N) gemm_epilogue(p, m, n, acc[i][j]);
    }
}

// ---- attention, head_dim 64, one thread per query (forward, dq) or per key (dk, dv) -----------------------------------
// qkv[B*L, >= 3*H*64] (q | k | v), token row = b*L + l.  lse[B*H*L] = rowmax + log(rowsum) of the SCALED scores (natural
// log — this buffer is private to the fp32 kernels).
constexpr int kAttnChunk = 32;

__device__ __forceinline__ void attn_stage(float (*dst)[64], const float* src, long long ld, int n_valid) {
  // 32 rows x 64 floats by 128 threads (float4 each, 4 passes); rows beyond n_valid are zero
  for (int e = threadIdx.x; e < kAttnChunk * 16; e += 128) {
    const int r = e >> 4, c = (e & 15) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < n_valid) v = *reinterpret_cast<const float4*>(src + (long long)r * ld + c);
    *reinterpret_cast<float4*>(&dst[r][c]) = v;
  }
}
__device__ __forceinline__ float dot64(const float (&a)[64], const float* b) {
  float s = 0.f;
#pragma unroll
  for (int d = 0; d < 64; d += 4) {
    const float4 v = *reinterpret_cast<const float4*>(b + d);
    s = fmaf(a[d], v.x, s); s = fmaf(a[d + 1], v.y, s); s = fmaf(a[d + 2], v.z, s); s = fmaf(a[d + 3], v.w, s);
  }
  return s;
}
__device__ __forceinline__ void load64(float (&r)[64], const float* p, bool ok) {
#pragma unroll
  for (int d = 0; d < 64; d += 4) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ok) v = *reinterpret_cast<const float4*>(p + d);
    r[d] = v.x; r[d + 1] = v.y; r[d + 2] = v.z; r[d + 3] = v.w;
  }
}
__device__ __forceinline__ void store64(float* p, const float (&r)[64], float mult) {
#pragma unroll
  for (int d = 0; d < 64; d += 4) *reinterpret_cast<float4*>(p + d) = make_float4(r[d] * mult, r[d + 1] * mult, r[d + 2] * mult, r[d + 3] * mult);
}

__global__ void __launch_bounds__(128) attn_fwd_kernel(const float* __restrict__ qkv, long long ld, float* __restrict__ o,
                                                      long long ldo, float* __restrict__ lse, int B, int H, int L, float scale,
                                                      int causal = 0) {
  // causal != 0: query l attends to keys j <= l only (the additive -inf upper-triangular mask of the CLIP text tower,
  // FS/model.py:1303-1309 build_attention_mask)
  __shared__ __align__(16) float Ks[kAttnChunk][64];
  __shared__ __align__(16) float Vs[kAttnChunk][64];
  const int bh = blockIdx.y, b = bh / H, h = bh % H, D = H * 64;
  const int l = blockIdx.x * 128 + threadIdx.x;
  const bool ok = l < L;
  const float* base = qkv + (long long)b * L * ld + h * 64;
  float q[64], acc[64];
  load64(q, base + (long long)l * ld, ok);
#pragma unroll
  for (int d = 0; d < 64; ++d) { q[d] *= scale; acc[d] = 0.f; }
  float m = -INFINITY, sum = 0.f;
  for (int j0 = 0; j0 < L; j0 += kAttnChunk) {
    const int nk = min(kAttnChunk, L - j0);
    __syncthreads();
    attn_stage(Ks, base + D + (long long)j0 * ld, ld, nk);
    attn_stage(Vs, base + 2 * D + (long long)j0 * ld, ld, nk);
    __syncthreads();
    float s[kAttnChunk];
    float cmax = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < kAttnChunk; ++jj) {
      s[jj] = (jj < nk && (!causal || j0 + jj <= l)) ? dot64(q, Ks[jj]) : -INFINITY;
      cmax = fmaxf(cmax, s[jj]);
    }
    const float mn = fmaxf(m, cmax);
    const float corr = expf(m - mn);
    sum *= corr;
#pragma unroll
    for (int d = 0; d < 64; ++d) acc[d] *= corr;
#pragma unroll
    for (int jj = 0; jj < kAttnChunk; ++jj) {
      const float pj = expf(s[jj] - mn);       // exp(-inf) = 0 for the padded keys
      sum += pj;
#pragma unroll
      for (int d = 0; d < 64; ++d) acc[d] = fmaf(pj, Vs[jj][d], acc[d]);
    }
    m = mn;
  }
  if (ok) {
    store64(o + ((long long)b * L + l) * ldo + h * 64, acc, 1.0f / sum);
    lse[(long long)bh * L + l] = m + logf(sum);
  }
}

// dq[i] = scale * sum_j ds_ij k_j,  ds_ij = p_ij (do_i . v_j - delta_i),  delta_i = do_i . o_i   (also written out)
__global__ void __launch_bounds__(128) attn_bwd_dq_kernel(const float* __restrict__ qkv, long long ld, const float* __restrict__ o,
                                                         long long ldo, const float* __restrict__ d_o, long long lddo,
                                                         const float* __restrict__ lse, float* __restrict__ delta,
                                                         float* __restrict__ dqkv, long long ldd, int B, int H, int L, float scale) {
  __shared__ __align__(16) float Ks[kAttnChunk][64];
  __shared__ __align__(16) float Vs[kAttnChunk][64];
  const int bh = blockIdx.y, b = bh / H, h = bh % H, D = H * 64;
  const int l = blockIdx.x * 128 + threadIdx.x;
  const bool ok = l < L;
  const float* base = qkv + (long long)b * L * ld + h * 64;
  const long long row = (long long)b * L + l;
  float q[64], g[64], dq[64];
  load64(q, base + (long long)l * ld, ok);
  load64(g, d_o + row * lddo + h * 64, ok);
  float dl = 0.f;
  {
    float ov[64];
    load64(ov, o + row * ldo + h * 64, ok);
#pragma unroll
    for (int d = 0; d < 64; ++d) dl = fmaf(g[d], ov[d], dl);
  }
#pragma unroll
  for (int d = 0; d < 64; ++d) { q[d] *= scale; dq[d] = 0.f; }
  const float ls = ok ? lse[(long long)bh * L + l] : 0.f;
  for (int j0 = 0; j0 < L; j0 += kAttnChunk) {
    const int nk = min(kAttnChunk, L - j0);
    __syncthreads();
    attn_stage(Ks, base + D + (long long)j0 * ld, ld, nk);
    attn_stage(Vs, base + 2 * D + (long long)j0 * ld, ld, nk);
    __syncthreads();
    for (int jj = 0; jj < nk; ++jj) {
      const float p = expf(dot64(q, Ks[jj]) - ls);
      const float ds = p * (dot64(g, Vs[jj]) - dl);
#pragma unroll
      for (int d = 0; d < 64; ++d) dq[d] = fmaf(ds, Ks[jj][d], dq[d]);
    }
  }
  if (ok) {
    store64(dqkv + row * ldd + h * 64, dq, scale);
    delta[(long long)bh * L + l] = dl;
  }
}

// one thread per key j.  WHICH = 0: dk_j = scale * sum_i ds_ij q_i      WHICH = 1: dv_j = sum_i p_ij do_i
template <int WHICH>
__global__ void __launch_bounds__(128) attn_bwd_dkv_kernel(const float* __restrict__ qkv, long long ld, const float* __restrict__ d_o,
                                                          long long lddo, const float* __restrict__ lse,
                                                          const float* __restrict__ delta, float* __restrict__ dqkv,
                                                          long long ldd, int B, int H, int L, float scale) {
  __shared__ __align__(16) float Qs[kAttnChunk][64];
  __shared__ __align__(16) float Gs[kAttnChunk][64];
  __shared__ float Ls[kAttnChunk], Ds[kAttnChunk];
  const int bh = blockIdx.y, b = bh / H, h = bh % H, D = H * 64;
  const int j = blockIdx.x * 128 + threadIdx.x;
  const bool ok = j < L;
  const float* base = qkv + (long long)b * L * ld + h * 64;
  const long long row = (long long)b * L + j;
  float k[64], v[64], acc[64];
  load64(k, base + D + (long long)j * ld, ok);
  if (WHICH == 0) load64(v, base + 2 * D + (long long)j * ld, ok);
#pragma unroll
  for (int d = 0; d < 64; ++d) { k[d] *= scale; acc[d] = 0.f; }
  for (int i0 = 0; i0 < L; i0 += kAttnChunk) {
    const int nq = min(kAttnChunk, L - i0);
    __syncthreads();
    attn_stage(Qs, base + (long long)i0 * ld, ld, nq);
    attn_stage(Gs, d_o + ((long long)b * L + i0) * lddo + h * 64, lddo, nq);
    if (threadIdx.x < kAttnChunk) {
      const bool in = (int)threadIdx.x < nq;
      Ls[threadIdx.x] = in ? lse[(long long)bh * L + i0 + threadIdx.x] : 0.f;
      Ds[threadIdx.x] = in ? delta[(long long)bh * L + i0 + threadIdx.x] : 0.f;
    }
    __syncthreads();
    for (int ii = 0; ii < nq; ++ii) {
      const float p = expf(dot64(k, Qs[ii]) - Ls[ii]);
      if (WHICH == 0) {
        const float ds = p * (dot64(v, Gs[ii]) - Ds[ii]);
#pragma unroll
        for (int d = 0; d < 64; ++d) acc[d] = fmaf(ds, Qs[ii][d], acc[d]);
      } else {
#pragma unroll
        for (int d = 0; d < 64; ++d) acc[d] = fmaf(p, Gs[ii][d], acc[d]);
      }
    }
  }
  if (ok) store64(dqkv + row * ldd + (WHICH == 0 ? D : 2 * D) + h * 64, acc, WHICH == 0 ? scale : 1.0f);
}

// ---- CLIP-dialect q/v update placement, fp32 (same chunk map as clip_delta_sm100.cuh) --------------------------------
template <bool BWD>
__global__ void __launch_bounds__(256) clip_delta_kernel(float* __restrict__ qkv, long long ldqkv, float* __restrict__ delta,
                                                        long long ldd, int B, int H, int L, float mult_q, float mult_v) {
  const long long chunks = (long long)B * H * L;
  const long long total = chunks * 2 * 16;             // x {q, v} x 16 float4
  const int D = H * 64;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int e4 = (int)(idx & 15);
    long long g = idx >> 4;
    const int which = g >= chunks ? 1 : 0;
    g -= which ? chunks : 0;
    const int l = (int)(g % L);
    const long long bh = g / L;
    const int h = (int)(bh % H), b = (int)(bh / H);
    const int slot = (int)(g % H);
    const long long r = g / H;
    const int b_s = (int)(r % B), l_s = (int)(r / B);
    float4* pd = reinterpret_cast<float4*>(qkv + ((long long)b * L + l) * ldqkv + (which ? 2 * D : 0) + h * 64) + e4;
    float4* ps = reinterpret_cast<float4*>(delta + ((long long)b_s * L + l_s) * ldd + which * D + slot * 64) + e4;
    const float mult = which ? mult_v : mult_q;
    if (BWD) {
      const float4 v = *pd;
      *ps = make_float4(v.x * mult, v.y * mult, v.z * mult, v.w * mult);
    } else {
      float4 v = *pd;
      const float4 s = *ps;
      v.x = fmaf(s.x, mult, v.x); v.y = fmaf(s.y, mult, v.y); v.z = fmaf(s.z, mult, v.z); v.w = fmaf(s.w, mult, v.w);
      *pd = v;
    }
  }
}

// ---- patch unfold, token assembly, vector add ------------------------------------------------------------------------
__global__ void im2col_kernel(const float* __restrict__ img, float* __restrict__ out, long long ldo, int B, int HW, int p) {
  const int gp = HW / p, K = 3 * p * p;
  const long long total = (long long)B * gp * gp * K;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(e % K);
    const long long row = e / K;
    const int c = k / (p * p), i = (k / p) % p, j = k % p;
    const int px = (int)(row % gp), py = (int)((row / gp) % gp);
    const int b = (int)(row / (gp * gp));
    out[row * ldo + k] = img[(((long long)b * 3 + c) * HW + (py * p + i)) * HW + px * p + j];
  }
}
__global__ void assemble_kernel(const float* __restrict__ pe, const float* __restrict__ cls, const float* __restrict__ pos,
                                float* __restrict__ x, int B, int L, int D) {
  const long long total = (long long)B * L * D;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(e % D);
    const long long row = e / D;
    const int l = (int)(row % L);
    const long long b = row / L;
    x[e] = (l == 0 ? cls[d] : pe[(b * (L - 1) + (l - 1)) * D + d]) + pos[(long long)l * D + d];
  }
}
__global__ void add_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, long long n) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) out[e] = a[e] + b[e];
}

}  // namespace f32
}  // namespace peft
