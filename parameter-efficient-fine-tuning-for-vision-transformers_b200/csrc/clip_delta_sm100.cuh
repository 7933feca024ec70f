// CLIP-dialect q/v delta placement (SURVEY.md §8a quirks Q1 + Q2).
//
// The reference adds the adapter update to the per-head q / v tensors through a RAW reshape of an [L, B, C] tensor to
// [B*H, L, hd] (FS/model.py:981-986, FS/lora_model.py:730-735) while q itself is view(L, B*H, hd).transpose(0, 1)
// (FS/model.py:907), so the 64-wide chunk that lands on (b, h, l) is the one whose flat chunk index in the [L, B, C]
// tensor equals g = (b*H + h)*L + l:   (r, slot) = divmod(g, H),  (l_s, b_s) = divmod(r, B),
//       q[b, h, l, :] += delta_q[l_s, b_s, slot*64 : slot*64+64]        (same for v).
// q has already been divided by sqrt(hd) at that point (Q2, FS/model.py:967); the attention kernels apply `scale` to
// the scores instead, so the q chunk is multiplied by mult_q = 1/scale here.
//
// Buffers use this library's token order (row = b*L + l):  qkv [B*L, >= 3*H*64] (q | k | v), delta [B*L, 2*H*64]
// (delta_q | delta_v).  The map chunk -> chunk is a bijection, so forward is a gather-add and backward a pure gather:
//   FWD:  qkv[dst] += mult * delta[src]          BWD:  ddelta[src] = mult * dqkv[dst]
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "pdl_sm100.cuh"

namespace peft {

template <bool BWD>
__global__ void __launch_bounds__(256) clip_delta_kernel(__nv_bfloat16* __restrict__ qkv, long long ldqkv,
                                                        __nv_bfloat16* __restrict__ delta, long long ldd, int B, int H,
                                                        int L, float mult_q, float mult_v) {
  pdl_trigger();                     // PDL (common.cuh)
  pdl_wait();
  const long long chunks = (long long)B * H * L;      // 64-element chunks per tensor (q or v)
  const long long total = chunks * 2 * 8;             // x {q, v} x 8 vectors of 8 bf16
  const int D = H * 64;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int e8 = (int)(idx & 7);
    long long g = idx >> 3;
    const int which = g >= chunks ? 1 : 0;
    g -= which ? chunks : 0;
    // destination (b, h, l)
    const int l = (int)(g % L);
    const long long bh = g / L;
    const int h = (int)(bh % H);
    const int b = (int)(bh / H);
    // source (l_s, b_s, slot)
    const int slot = (int)(g % H);
    const long long r = g / H;
    const int b_s = (int)(r % B);
    const int l_s = (int)(r / B);
    __nv_bfloat16* pd = qkv + ((long long)b * L + l) * ldqkv + (which ? 2 * D : 0) + h * 64 + e8 * 8;
    __nv_bfloat16* ps = delta + ((long long)b_s * L + l_s) * ldd + which * D + slot * 64 + e8 * 8;
    const float mult = which ? mult_v : mult_q;
    uint4 vs, vd;
    if (BWD) {
      vd = *reinterpret_cast<const uint4*>(pd);
      const __nv_bfloat162* a = reinterpret_cast<const __nv_bfloat162*>(&vd);
      __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&vs);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float2 f = __bfloat1622float2(a[k]);
        o[k] = __floats2bfloat162_rn(f.x * mult, f.y * mult);
      }
      *reinterpret_cast<uint4*>(ps) = vs;
    } else {
      vd = *reinterpret_cast<const uint4*>(pd);
      vs = *reinterpret_cast<const uint4*>(ps);
      const __nv_bfloat162* a = reinterpret_cast<const __nv_bfloat162*>(&vd);
      const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(&vs);
      uint4 out;
      __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&out);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float2 f = __bfloat1622float2(a[k]);
        float2 dd = __bfloat1622float2(d[k]);
        o[k] = __floats2bfloat162_rn(fmaf(dd.x, mult, f.x), fmaf(dd.y, mult, f.y));
      }
      *reinterpret_cast<uint4*>(pd) = out;
    }
  }
}

// out[c, r] = in[r, c] for a bf16 matrix [rows, cols] (pitch ldin) -> [cols, rows] (pitch ldout).  Used to bring the two
// operands of a weight-gradient product dW = dY^T X (contraction over the token dimension) into the K-major form the
// tcgen05 GEMM consumes (full fine-tuning / intrinsic-dimension mode only, SURVEY R12: "backward needs full dW").
__global__ void __launch_bounds__(256) transpose_bf16_kernel(const __nv_bfloat16* __restrict__ in, long long ldin, int rows,
                                                            int cols, __nv_bfloat16* __restrict__ out, long long ldout) {
  __shared__ __nv_bfloat16 tile[64][66];
  pdl_trigger();                     // PDL (common.cuh)
  pdl_wait();
  const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
  const int lane = threadIdx.x & 31, wy = threadIdx.x >> 5;
  // load: 8 warps x 8 passes, each warp reads one 64-element row as 32 x bf16x2
#pragma unroll
  for (int p = 0; p < 8; ++p) {
    const int r = p * 8 + wy;
    const int gr = r0 + r, gc = c0 + lane * 2;
    __nv_bfloat16 a = __float2bfloat16(0.f), b = a;
    if (gr < rows) {
      const __nv_bfloat16* src = in + (long long)gr * ldin + gc;
      if (gc + 1 < cols) {
        const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(src);
        a = v.x; b = v.y;
      } else if (gc < cols) {
        a = src[0];
      }
    }
    tile[r][lane * 2] = a;
    tile[r][lane * 2 + 1] = b;
  }
  __syncthreads();
  // store: output row = c, 64 consecutive r per row, lane writes r = 2*lane, 2*lane+1
#pragma unroll
  for (int p = 0; p < 8; ++p) {
    const int c = p * 8 + wy;
    const int gc = c0 + c, gr = r0 + lane * 2;
    if (gc < cols) {
      __nv_bfloat16* dst = out + (long long)gc * ldout + gr;
      if (gr + 1 < rows) {
        __nv_bfloat162 v;
        v.x = tile[lane * 2][c];
        v.y = tile[lane * 2 + 1][c];
        *reinterpret_cast<__nv_bfloat162*>(dst) = v;
      } else if (gr < rows) {
        dst[0] = tile[lane * 2][c];
      }
    }
  }
}

// 16-byte-aligned fast path: no shared memory.  Each thread moves one 8 x 8 bf16 block: 8 row vectors (16 B, 8 columns)
// in, an in-register transpose with byte permutes, 8 column vectors (16 B, 8 rows) out.  Lanes are laid out 4 (column
// groups) x 8 (row groups), so a warp reads 64-byte and writes 128-byte contiguous runs: whole 32-byte sectors both ways.
// CTA = 8 warps side by side: 64 rows x 256 columns.  Blocks crossing the matrix edge take the guarded scalar path.
// COLSUM: also colsum[c] += sum_r in[r, c] (fp32): the bias gradient of a linear layer is the column sum of the same dY
// that is being transposed for the weight-gradient product, so it rides along (8 lanes share a column group: three
// shuffles, then one global atomic per column per warp) instead of costing another pass over dY.
// STORE = false: column sums only (nothing is written to `out`): the bias gradient when dW comes from peft_gemm_tn_bf16.
template <bool COLSUM, bool STORE = true>
__global__ void __launch_bounds__(256) transpose_bf16_vec_kernel(const __nv_bfloat16* __restrict__ in, long long ldin, int rows,
                                                                int cols, __nv_bfloat16* __restrict__ out, long long ldout,
                                                                float* __restrict__ colsum) {
  pdl_trigger();                     // PDL (common.cuh)
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cg = lane & 3, rg = lane >> 2;
  const int r0 = blockIdx.y * 64 + rg * 8;
  const int c0 = blockIdx.x * 256 + warp * 32 + cg * 8;
  float cs[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) cs[c] = 0.f;
  if (r0 >= rows || c0 >= cols) {
    if (!COLSUM) return;
  } else if (r0 + 8 <= rows && c0 + 8 <= cols) {
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = *reinterpret_cast<const uint4*>(in + (long long)(r0 + j) * ldin + c0);
    if (COLSUM) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[j]);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 f = __bfloat1622float2(h[q]);
          cs[2 * q] += f.x;
          cs[2 * q + 1] += f.y;
        }
      }
    }
#pragma unroll
    for (int c = 0; c < (STORE ? 8 : 0); ++c) {
      const uint32_t sel = (c & 1) ? 0x7632u : 0x5410u;      // high / low bf16 of both source registers
      uint32_t w[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t a = reinterpret_cast<const uint32_t*>(&v[2 * k])[c >> 1];
        const uint32_t b = reinterpret_cast<const uint32_t*>(&v[2 * k + 1])[c >> 1];
        w[k] = __byte_perm(a, b, sel);
      }
      *reinterpret_cast<uint4*>(out + (long long)(c0 + c) * ldout + r0) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  } else {
    for (int j = 0; j < 8 && r0 + j < rows; ++j)
      for (int c = 0; c < 8 && c0 + c < cols; ++c) {
        const __nv_bfloat16 e = in[(long long)(r0 + j) * ldin + c0 + c];
        if (STORE) out[(long long)(c0 + c) * ldout + r0 + j] = e;
        if (COLSUM) cs[c] += __bfloat162float(e);
      }
  }
  if (COLSUM) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {                       // lanes with equal (lane & 3) hold the same 8 columns
      cs[c] += __shfl_xor_sync(0xffffffffu, cs[c], 4);
      cs[c] += __shfl_xor_sync(0xffffffffu, cs[c], 8);
      cs[c] += __shfl_xor_sync(0xffffffffu, cs[c], 16);
    }
    if (rg == 0) {
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if (c0 + c < cols) atomicAdd(colsum + c0 + c, cs[c]);
    }
  }
}

// colsum[c] += sum_r in[r, c]: streaming column sums (bias gradients).  Same lane layout as the transpose (4 column groups
// x 8 row groups per warp, 16-byte loads, 64-byte runs per row); each CTA walks several 64-row blocks and accumulates in
// registers, so a column receives gridDim.y atomics in total instead of one per 64 rows (the one-block-per-CTA version
// spent more time serialising same-address atomics in L2 than streaming: 23 us for [12800, 3072]).
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ in, long long ldin, int rows, int cols,
                                                         float* __restrict__ colsum) {
  pdl_trigger();                     // PDL (common.cuh)
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cg = lane & 3, rg = lane >> 2;
  const int c0 = blockIdx.x * 256 + warp * 32 + cg * 8;
  float cs[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) cs[c] = 0.f;
  if (c0 < cols) {
    const bool full = c0 + 8 <= cols;
    for (int rb = blockIdx.y; rb * 64 < rows; rb += gridDim.y) {
      const int r0 = rb * 64 + rg * 8;
      if (full) {
        uint4 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
          v[j] = (r0 + j < rows) ? *reinterpret_cast<const uint4*>(in + (long long)(r0 + j) * ldin + c0) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[j]);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float2 f = __bfloat1622float2(h[q]);
            cs[2 * q] += f.x;
            cs[2 * q + 1] += f.y;
          }
        }
      } else {
        for (int j = 0; j < 8 && r0 + j < rows; ++j)
          for (int c = 0; c < 8 && c0 + c < cols; ++c) cs[c] += __bfloat162float(in[(long long)(r0 + j) * ldin + c0 + c]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 8; ++c) {                         // lanes with equal (lane & 3) hold the same 8 columns
    cs[c] += __shfl_xor_sync(0xffffffffu, cs[c], 4);
    cs[c] += __shfl_xor_sync(0xffffffffu, cs[c], 8);
    cs[c] += __shfl_xor_sync(0xffffffffu, cs[c], 16);
  }
  if (rg == 0) {
#pragma unroll
    for (int c = 0; c < 8; ++c)
      if (c0 + c < cols) atomicAdd(colsum + c0 + c, cs[c]);
  }
}

}  // namespace peft

// ---- large-n KAdaptation "fold" (the committed phm_dim = n = D, K = P = 1) --------------------------
// H_p[j] = sum_i s_p[i] * A[i, j]   for p in {q, v}, j in [0, n*n):  the weighted sum of the n rule matrices, i.e.
// kronecker_product_einsum_batched(phm_rule, W).sum(0) of FS/kadaptation.py:81-115 when every W_i is a 1x1 block.
// HBM-bound: A (= phm_rule, n^3 fp32 = 1.81 GB at n = 768) is read exactly once for BOTH projections.
namespace peft {
constexpr int kFoldThreads = 256;
constexpr int kFoldMaxN = 1024;

__global__ void __launch_bounds__(kFoldThreads) phm_fold_fwd_kernel(const float* __restrict__ A, const float* __restrict__ s,
                                                                    float* __restrict__ H, int n, long long nn) {
  __shared__ float sq[kFoldMaxN], sv[kFoldMaxN];
  for (int i = threadIdx.x; i < n; i += kFoldThreads) { sq[i] = s[i]; sv[i] = s[n + i]; }
  __syncthreads();
  const long long j = ((long long)blockIdx.x * kFoldThreads + threadIdx.x) * 4;
  if (j >= nn) return;
  const float4* a = reinterpret_cast<const float4*>(A + j);
  const long long stride = nn / 4;
  float4 hq = make_float4(0.f, 0.f, 0.f, 0.f), hv = hq;
#pragma unroll 8
  for (int i = 0; i < n; ++i) {
    const float4 v = __ldcs(a + (long long)i * stride);         // streamed once: do not keep in L2
    const float wq = sq[i], wv = sv[i];
    hq.x = fmaf(wq, v.x, hq.x); hq.y = fmaf(wq, v.y, hq.y); hq.z = fmaf(wq, v.z, hq.z); hq.w = fmaf(wq, v.w, hq.w);
    hv.x = fmaf(wv, v.x, hv.x); hv.y = fmaf(wv, v.y, hv.y); hv.z = fmaf(wv, v.z, hv.z); hv.w = fmaf(wv, v.w, hv.w);
  }
  *reinterpret_cast<float4*>(H + j) = hq;
  *reinterpret_cast<float4*>(H + nn + j) = hv;
}

// ds_p[i] += sum_j dH_p[j] * A[i, j].  A CTA takes kFoldRows rule matrices and one of kFoldSplit slices of j, so every dH
// vector it pulls from L2 is used kFoldRows times (dH is 2 x n^2 fp32 = 4.7 MB: L2-resident) while A is streamed exactly
// once from HBM; partial sums go to ds with one atomic per (row, projection) per CTA (ds is zeroed by the host wrapper).
constexpr int kFoldRows = 4;
constexpr int kFoldSplit = 4;
__global__ void __launch_bounds__(kFoldThreads) phm_fold_bwd_kernel(const float* __restrict__ A, const float* __restrict__ dH,
                                                                    float* __restrict__ ds, int n, long long nn) {
  const int i0 = blockIdx.x * kFoldRows;
  const long long n4 = nn / 4;
  const long long per = (n4 + kFoldSplit - 1) / kFoldSplit;
  const long long k0 = blockIdx.y * per, k1 = (k0 + per < n4) ? k0 + per : n4;
  const float4* gq = reinterpret_cast<const float4*>(dH);
  const float4* gv = reinterpret_cast<const float4*>(dH + nn);
  const float4* a[kFoldRows];
#pragma unroll
  for (int r = 0; r < kFoldRows; ++r) a[r] = reinterpret_cast<const float4*>(A + (long long)((i0 + r) < n ? (i0 + r) : (n - 1)) * nn);
  float aq[kFoldRows], av[kFoldRows];
#pragma unroll
  for (int r = 0; r < kFoldRows; ++r) { aq[r] = 0.f; av[r] = 0.f; }
#pragma unroll 2
  for (long long k = k0 + threadIdx.x; k < k1; k += kFoldThreads) {
    const float4 q = __ldg(gq + k), w = __ldg(gv + k);
#pragma unroll
    for (int r = 0; r < kFoldRows; ++r) {
      const float4 v = __ldcs(a[r] + k);
      aq[r] = fmaf(v.x, q.x, fmaf(v.y, q.y, fmaf(v.z, q.z, fmaf(v.w, q.w, aq[r]))));
      av[r] = fmaf(v.x, w.x, fmaf(v.y, w.y, fmaf(v.z, w.z, fmaf(v.w, w.w, av[r]))));
    }
  }
  __shared__ float red[2 * kFoldRows][kFoldThreads / 32];
#pragma unroll
  for (int r = 0; r < kFoldRows; ++r) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      aq[r] += __shfl_xor_sync(0xffffffffu, aq[r], o);
      av[r] += __shfl_xor_sync(0xffffffffu, av[r], o);
    }
    if ((threadIdx.x & 31) == 0) { red[2 * r][threadIdx.x >> 5] = aq[r]; red[2 * r + 1][threadIdx.x >> 5] = av[r]; }
  }
  __syncthreads();
  if (threadIdx.x < 2 * kFoldRows) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < kFoldThreads / 32; ++k) t += red[threadIdx.x][k];
    const int r = threadIdx.x >> 1, p = threadIdx.x & 1;
    if (i0 + r < n) atomicAdd(&ds[p * n + i0 + r], t);
  }
}
}  // namespace peft
