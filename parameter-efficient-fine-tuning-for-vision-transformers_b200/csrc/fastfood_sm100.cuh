// Fastfood intrinsic-dimension projection: theta_p = theta0_p + c_p * [H G Pi H B pad(V)][:DD_p]
// for every wrapped parameter tensor p, and its transpose dV = sum_p c_p * B_p .* [H Pi^T G H pad(dtheta_p)][:d].
// replaces: FL/intrinsic/fastfood.py:61-77 (FastfoodWrap.forward), :81-121 (fast_walsh_hadamard_torched),
// :151-187 (fastfood_torched) and their autograd backward.
//
// Structure exploited (nothing here changes the arithmetic, only the order of the +/- sums):
//   * pad(V) has its non-zeros in the first d <= wn entries (wn = next power of two), and H_LL = H_{LL/wn} (x) H_wn in
//     natural (Sylvester) order, so  H_LL (B .* pad(V)) = tile(w, LL/wn)  with  w = H_wn (B[:wn] .* pad_wn(V)):
//     the first transform is ONE wn-point FWHT per tensor and the permutation only needs  Pi[j] mod wn  (uint16).
//   * transposed: [H_LL y][:wn] = H_wn fold(y),  fold(y)[k] = sum_{j = k mod wn} y[j], so the backward scatter through
//     Pi^T becomes a wn-bin accumulation (shared-memory atomics, then one global atomic per bin per CTA).
//   * the LL-point FWHT (LL up to 2^22 for fc1/fc2/conv1) is split into a strided "high" pass over bits [13, ll) and a
//     contiguous "low" pass over bits [0, 13), each an 8192-element shared-memory tile (512 threads x 16 registers,
//     radix-16 register butterflies, XOR-swizzled transposes).  Forward runs high -> low so that the low pass touches
//     only the blocks below DD; backward runs low (non-zero blocks only) -> high.
// HBM traffic per step: forward 10 B per LL element + 12 B per DD element, backward 8 B + (6+4) B; the single-pass
// algorithmic minimum is 6 B per LL element + 8 B (fwd) / 4 B (bwd) per DD element.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/peft_b200.h"

namespace peft {

constexpr int kFfThreads = 512;
constexpr int kFfTile = 8192;      // elements per tile = 2^13
constexpr int kFfTileBits = 13;
constexpr int kFfMaxWn = 4096;

__device__ __forceinline__ int ff_swz(int idx) { return idx ^ (((idx >> 5) & 15) << 1); }

// in-register butterflies on register-index bits 0..3 <-> element bits base..base+3, applied only to bits in [lc, hc)
__device__ __forceinline__ void ff_butterfly16(float (&x)[16], int base, int lc, int hc) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int bit = base + k;
    if (bit >= lc && bit < hc) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        if ((i & (1 << k)) == 0) {
          const float a = x[i], b = x[i | (1 << k)];
          x[i] = a + b;
          x[i | (1 << k)] = a - b;
        }
      }
    }
  }
}

// Unnormalised Walsh-Hadamard butterflies over element bits [lc, hc) of an 8192-element tile.  On entry and exit thread
// `tid` holds elements e = i*512 + tid (i = 0..15).  `s` is 8192 floats of shared memory (clobbered).
__device__ __forceinline__ void ff_tile_fwht(float (&x)[16], float* s, int lc, int hc) {
  const int tid = threadIdx.x;
  ff_butterfly16(x, 9, lc, hc);                       // bits 9..12 live in the register index
  if (lc >= 9 || hc <= 0) return;
#pragma unroll
  for (int i = 0; i < 16; ++i) s[ff_swz(i * 512 + tid)] = x[i];
  __syncthreads();
  if (hc > 5) {                                       // bits 5..8: idx = hi*512 + i*32 + lane
    const int lane = tid & 31, hi = tid >> 5;
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = s[ff_swz(hi * 512 + i * 32 + lane)];
    ff_butterfly16(x, 5, lc, hc);
#pragma unroll
    for (int i = 0; i < 16; ++i) s[ff_swz(hi * 512 + i * 32 + lane)] = x[i];
    __syncthreads();
  }
  if (lc < 5) {                                       // bits 1..4 in registers, bit 0 across the lane pair
    const int c0 = tid & 1, q = tid >> 1;
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = s[ff_swz(q * 32 + i * 2 + c0)];
    ff_butterfly16(x, 1, lc, hc);
    if (lc == 0) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float o = __shfl_xor_sync(0xffffffffu, x[i], 1);
        x[i] = c0 ? (o - x[i]) : (x[i] + o);
      }
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) s[ff_swz(q * 32 + i * 2 + c0)] = x[i];
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = s[ff_swz(i * 512 + tid)];
}

struct FfArgs {
  const peft_ff_tensor* tensors;
  const int2* items;            // (tensor, tile)
  const uint16_t* pi_lo;        // Pi[j] mod min(LL, wn), concatenated over tensors (off_ll)
  const float* gg;              // G, same layout
  const float* wv;              // [T, wn]: forward w = H (B .* V); unused in backward
  float* u;                     // [T, wn]: backward bins
  float* scratch;               // [sum LL] fp32 intermediate between the two passes
  const float* theta0;          // [sum DD (padded offsets)]
  float* theta;                 // forward output
  const float* dtheta;          // backward input
  int wn;
};

// ---- forward, high bits: x[j] = w[Pi[j] mod wn] * G[j] generated on the fly, butterflies over bits [13, ll) -----------
__global__ void __launch_bounds__(kFfThreads, 2) ff_fwd_high_kernel(FfArgs a) {
  extern __shared__ float ff_smem[];
  float* s = ff_smem;
  float* w = ff_smem + kFfTile;
  const int2 it = a.items[blockIdx.x];
  const peft_ff_tensor t = a.tensors[it.x];
  const int tid = threadIdx.x;
  for (int k = tid; k < a.wn; k += kFfThreads) w[k] = a.wv[(long long)it.x * a.wn + k];
  __syncthreads();
  const int lc = kFfTileBits - (t.ll - kFfTileBits);     // log2(columns per tile)
  const int cmask = (1 << lc) - 1;
  const long long base = t.off_ll + ((long long)it.y << lc);
  float x[16];
  auto jof = [&](int i) {
    const int e = i * 512 + tid;
    return base + ((long long)(e >> lc) << kFfTileBits) + (e & cmask);
  };
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const long long j = jof(i);
    x[i] = w[a.pi_lo[j]] * a.gg[j];
  }
  ff_tile_fwht(x, s, lc, kFfTileBits);
#pragma unroll
  for (int i = 0; i < 16; ++i) a.scratch[jof(i)] = x[i];
}

// ---- forward, low bits: butterflies over bits [0, min(ll, 13)), theta = theta0 + coef * x for j < DD ------------------
__global__ void __launch_bounds__(kFfThreads, 2) ff_fwd_low_kernel(FfArgs a) {
  extern __shared__ float ff_smem[];
  float* s = ff_smem;
  float* w = ff_smem + kFfTile;
  const int2 it = a.items[blockIdx.x];
  const peft_ff_tensor t = a.tensors[it.x];
  const int tid = threadIdx.x;
  const bool gen = t.ll <= kFfTileBits;                 // single-pass tensor: generate the input here
  const long long LL = 1ll << t.ll;
  const long long blk = (long long)it.y << kFfTileBits;
  float x[16];
  if (gen) {
    for (int k = tid; k < a.wn; k += kFfThreads) w[k] = a.wv[(long long)it.x * a.wn + k];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const long long j = blk + i * 512 + tid;
      x[i] = j < LL ? w[a.pi_lo[t.off_ll + j]] * a.gg[t.off_ll + j] : 0.f;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = a.scratch[t.off_ll + blk + i * 512 + tid];
  }
  ff_tile_fwht(x, s, 0, t.ll < kFfTileBits ? t.ll : kFfTileBits);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const long long j = blk + i * 512 + tid;
    if (j < t.DD) a.theta[t.off_dd + j] = a.theta0[t.off_dd + j] + t.coef * x[i];
  }
}

// ---- backward, low bits: x = pad(dtheta), butterflies over bits [0, min(ll,13)); single-pass tensors bin here ---------
__global__ void __launch_bounds__(kFfThreads, 2) ff_bwd_low_kernel(FfArgs a) {
  extern __shared__ float ff_smem[];
  float* s = ff_smem;
  float* bins = ff_smem + kFfTile;
  const int2 it = a.items[blockIdx.x];
  const peft_ff_tensor t = a.tensors[it.x];
  const int tid = threadIdx.x;
  const bool fin = t.ll <= kFfTileBits;
  const long long LL = 1ll << t.ll;
  const long long blk = (long long)it.y << kFfTileBits;
  float x[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const long long j = blk + i * 512 + tid;
    x[i] = j < t.DD ? a.dtheta[t.off_dd + j] : 0.f;
  }
  if (fin) {
    for (int k = tid; k < a.wn; k += kFfThreads) bins[k] = 0.f;
  }
  ff_tile_fwht(x, s, 0, t.ll < kFfTileBits ? t.ll : kFfTileBits);
  if (!fin) {
#pragma unroll
    for (int i = 0; i < 16; ++i) a.scratch[t.off_ll + blk + i * 512 + tid] = x[i];
    return;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const long long j = blk + i * 512 + tid;
    if (j < LL) atomicAdd(&bins[a.pi_lo[t.off_ll + j]], x[i] * a.gg[t.off_ll + j]);
  }
  __syncthreads();
  const int n1 = LL < a.wn ? (int)LL : a.wn;
  for (int k = tid; k < n1; k += kFfThreads) atomicAdd(&a.u[(long long)it.x * a.wn + k], bins[k]);
}

// ---- backward, high bits: butterflies over bits [13, ll), times G, folded through Pi^T into wn bins -------------------
__global__ void __launch_bounds__(kFfThreads, 2) ff_bwd_high_kernel(FfArgs a) {
  extern __shared__ float ff_smem[];
  float* s = ff_smem;
  float* bins = ff_smem + kFfTile;
  const int2 it = a.items[blockIdx.x];
  const peft_ff_tensor t = a.tensors[it.x];
  const int tid = threadIdx.x;
  for (int k = tid; k < a.wn; k += kFfThreads) bins[k] = 0.f;
  const int lc = kFfTileBits - (t.ll - kFfTileBits);
  const int cmask = (1 << lc) - 1;
  const long long base = t.off_ll + ((long long)it.y << lc);
  const int nblk = (t.DD + kFfTile - 1) >> kFfTileBits;   // rows (low blocks) that bwd_low wrote; the rest are zero
  float x[16];
  auto jof = [&](int i) {
    const int e = i * 512 + tid;
    return base + ((long long)(e >> lc) << kFfTileBits) + (e & cmask);
  };
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int row = (i * 512 + tid) >> lc;
    x[i] = row < nblk ? a.scratch[jof(i)] : 0.f;
  }
  ff_tile_fwht(x, s, lc, kFfTileBits);
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const long long j = jof(i);
    atomicAdd(&bins[a.pi_lo[j]], x[i] * a.gg[j]);
  }
  __syncthreads();
  for (int k = tid; k < a.wn; k += kFfThreads) atomicAdd(&a.u[(long long)it.x * a.wn + k], bins[k]);
}

// ---- the wn-point transforms at the V end (one CTA per tensor) -------------------------------------------------------
// MODE 0: w[t] = H_n1 (B_t[:n1] .* pad(V))          MODE 1: dV[k] += coef_t * B_t[k] * (H_n1 u[t])[k],  k < min(d, n1)
template <int MODE>
__global__ void __launch_bounds__(256) ff_small_kernel(const peft_ff_tensor* tensors, const float* bsign, const float* V,
                                                       float* wu, float* dV, int d, int wn) {
  extern __shared__ float ff_small_smem[];
  float* s = ff_small_smem;
  const int t = blockIdx.x;
  const peft_ff_tensor tt = tensors[t];
  const long long LL = 1ll << tt.ll;
  const int n1 = LL < wn ? (int)LL : wn;
  const float* bs = bsign + (long long)t * wn;
  float* row = wu + (long long)t * wn;
  for (int k = threadIdx.x; k < n1; k += blockDim.x) s[k] = MODE == 0 ? (k < d ? bs[k] * V[k] : 0.f) : row[k];
  __syncthreads();
  for (int h = 1; h < n1; h <<= 1) {
    for (int p = threadIdx.x; p < n1 / 2; p += blockDim.x) {
      const int i0 = ((p & ~(h - 1)) << 1) | (p & (h - 1));
      const float x0 = s[i0], x1 = s[i0 + h];
      s[i0] = x0 + x1;
      s[i0 + h] = x0 - x1;
    }
    __syncthreads();
  }
  if (MODE == 0) {
    for (int k = threadIdx.x; k < n1; k += blockDim.x) row[k] = s[k];
  } else {
    const int lim = d < n1 ? d : n1;
    for (int k = threadIdx.x; k < lim; k += blockDim.x) atomicAdd(&dV[k], tt.coef * bs[k] * s[k]);
  }
}

}  // namespace peft
