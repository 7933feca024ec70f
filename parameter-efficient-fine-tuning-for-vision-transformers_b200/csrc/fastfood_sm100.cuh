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
//   * both passes of one direction run in ONE launch: the work list is ordered tensor by tensor (forward: high tiles
//     then low blocks; backward: low blocks then high tiles) and the second pass of a tensor waits on a per-tensor
//     completion counter, so the fp32 intermediate of a tensor (<= 16 MB) is still in the 126 MB L2 when it is re-read
//     (read with ld.global.cg).  CTAs are dispatched in blockIdx order, and producers never wait, so the spin is safe.
// HBM traffic per step is then the algorithmic minimum: 6 B per LL element (uint16 Pi mod wn + fp32 G) plus 8 B (fwd:
// theta0 read, theta write) or 4 B (bwd: dtheta read) per DD element.
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
// Shared-memory layout: element idx lives at ff_swz(idx) = idx ^ (((idx >> 5) & 15) << 1); the three access patterns
// below are bank-conflict free under it, and each reduces to a per-thread constant plus a compile-time term.
__device__ __forceinline__ void ff_tile_fwht(float (&x)[16], float* s, int lc, int hc) {
  const int tid = threadIdx.x;
  ff_butterfly16(x, 9, lc, hc);                       // bits 9..12 live in the register index
  if (lc >= 9 || hc <= 0) return;
  float* sa = s + (tid ^ (((tid >> 5) & 15) << 1));   // ff_swz(i*512 + tid) = i*512 + (tid ^ ...)
#pragma unroll
  for (int i = 0; i < 16; ++i) sa[i * 512] = x[i];
  __syncthreads();
  if (hc > 5) {                                       // bits 5..8: idx = hi*512 + i*32 + lane, swz = idx ^ (i << 1)
    const int lane = tid & 31;
    float* sb = s + (tid >> 5) * 512;
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = sb[i * 32 + (lane ^ (i << 1))];
    ff_butterfly16(x, 5, lc, hc);
#pragma unroll
    for (int i = 0; i < 16; ++i) sb[i * 32 + (lane ^ (i << 1))] = x[i];
    __syncthreads();
  }
  if (lc < 5) {                                       // bits 1..4 in registers, bit 0 across the lane pair
    const int c0 = tid & 1, q = tid >> 1;             // idx = q*32 + i*2 + c0, swz = idx ^ ((q & 15) << 1)
    float* sc = s + q * 32 + c0;
    const int r2 = (q & 15) << 1;
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = sc[(i << 1) ^ r2];
    ff_butterfly16(x, 1, lc, hc);
    if (lc == 0) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float o = __shfl_xor_sync(0xffffffffu, x[i], 1);
        x[i] = c0 ? (o - x[i]) : (x[i] + o);
      }
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) sc[(i << 1) ^ r2] = x[i];
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = sa[i * 512];
}

struct FfArgs {
  const peft_ff_tensor* tensors;
  const int2* items;            // (tensor, tile | kind << 30): kind 0 = low block, 1 = high tile; ordered tensor by tensor
  const uint16_t* pi_lo;        // Pi[j] mod min(LL, wn), concatenated over tensors (off_ll)
  const float* gg;              // G, same layout
  const float* wv;              // [T, wn]: forward w = H (B .* V); unused in backward
  float* u;                     // [T, wn]: backward bins
  float* scratch;               // [sum LL] fp32 intermediate between the two passes
  const float* theta0;          // [sum DD (padded offsets)]
  float* theta;                 // forward output
  const float* dtheta;          // backward input
  int* counters;                // [T] completed first-pass CTAs per tensor (zeroed by the host wrapper)
  int wn;
};

// offset of element e = i*512 + tid of a high-pass tile inside its tensor: ((e >> lc) << 13) + (e & (2^lc - 1)), split
// into a per-thread constant and a term that only depends on (i, lc) and therefore lives on the uniform datapath
__device__ __forceinline__ int ff_hi_tconst(int tid, int lc) {
  return lc >= 9 ? tid : (((tid >> lc) << kFfTileBits) + (tid & ((1 << lc) - 1)));
}
__device__ __forceinline__ int ff_hi_ioff(int i, int lc) {
  return lc >= 9 ? (((i >> (lc - 9)) << kFfTileBits) + ((i & ((1 << (lc - 9)) - 1)) << 9)) : (i << (22 - lc));
}
__device__ __forceinline__ void ff_signal(int* counter) {       // all of this CTA's global writes -> visible, then count
  __syncthreads();                                               // bar.sync orders every thread's stores before thread 0's
  if (threadIdx.x == 0) {                                        // fence (cumulativity): one membar per CTA instead of 512
    __threadfence();
    atomicAdd(counter, 1);
  }
}
__device__ __forceinline__ void ff_wait(const int* counter, int target) {
  if (threadIdx.x == 0) {
    int v;
    do {
      asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
      if (v < target) __nanosleep(200);
    } while (v < target);
  }
  __syncthreads();
}

// ---- forward ---------------------------------------------------------------------------------------------------------
// high tile: x[j] = w[Pi[j] mod wn] * G[j] generated on the fly, butterflies over bits [13, ll), fp32 intermediate out
// low block: butterflies over bits [0, min(ll, 13)), theta = theta0 + coef * x for j < DD
__global__ void __launch_bounds__(kFfThreads, 2) ff_fwd_kernel(FfArgs a) {
  extern __shared__ float ff_smem[];
  float* s = ff_smem;
  float* w = ff_smem + kFfTile;
  const int2 it = a.items[blockIdx.x];
  const peft_ff_tensor t = a.tensors[it.x];
  const int tid = threadIdx.x;
  const int tile = it.y & 0x3fffffff;
  const bool high = (it.y >> 30) != 0;
  const uint16_t* pi = a.pi_lo + t.off_ll;
  const float* gg = a.gg + t.off_ll;
  float* scr = a.scratch + t.off_ll;
  float x[16];
  const float* wv = a.wv + (long long)it.x * a.wn;
  if (high) {
    // all global loads of the tile are issued before the first barrier so that their latencies overlap
    const int lc = 2 * kFfTileBits - t.ll;               // log2(columns per tile)
    const int tc = ff_hi_tconst(tid, lc) + (tile << lc);
    uint16_t pv[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int j = tc + ff_hi_ioff(i, lc);
      pv[i] = pi[j];
      x[i] = gg[j];
    }
    for (int k = tid; k < a.wn; k += kFfThreads) w[k] = wv[k];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] *= w[pv[i]];
    ff_tile_fwht(x, s, lc, kFfTileBits);
#pragma unroll
    for (int i = 0; i < 16; ++i) scr[tc + ff_hi_ioff(i, lc)] = x[i];
    ff_signal(a.counters + it.x);
    return;
  }
  const int j0 = (tile << kFfTileBits) + tid;
  const float* th0 = a.theta0 + t.off_dd;
  float* th = a.theta + t.off_dd;
  float t0v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {                         // theta0 is independent of everything else: fetch it first
    const int j = j0 + i * 512;
    t0v[i] = j < t.DD ? th0[j] : 0.f;
  }
  if (t.ll <= kFfTileBits) {                             // single-pass tensor: generate the input here
    const int LL = 1 << t.ll;
    uint16_t pv[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int j = j0 + i * 512;
      pv[i] = j < LL ? pi[j] : (uint16_t)0;
      x[i] = j < LL ? gg[j] : 0.f;
    }
    for (int k = tid; k < a.wn; k += kFfThreads) w[k] = wv[k];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] *= w[pv[i]];
  } else {
    ff_wait(a.counters + it.x, 1 << (t.ll - kFfTileBits));
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = __ldcg(scr + j0 + i * 512);
  }
  ff_tile_fwht(x, s, 0, t.ll < kFfTileBits ? t.ll : kFfTileBits);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int j = j0 + i * 512;
    if (j < t.DD) th[j] = fmaf(t.coef, x[i], t0v[i]);
  }
}

// ---- backward --------------------------------------------------------------------------------------------------------
// low block: x = pad(dtheta), butterflies over bits [0, min(ll,13)); single-pass tensors bin here, others write fp32
// high tile: butterflies over bits [13, ll), times G, folded through Pi^T into wn bins (shared atomics, then global)
__global__ void __launch_bounds__(kFfThreads, 3) ff_bwd_kernel(FfArgs a) {
  extern __shared__ float ff_smem[];
  float* s = ff_smem;
  float* bins = ff_smem + kFfTile;
  const int2 it = a.items[blockIdx.x];
  const peft_ff_tensor t = a.tensors[it.x];
  const int tid = threadIdx.x;
  const int tile = it.y & 0x3fffffff;
  const bool high = (it.y >> 30) != 0;
  const uint16_t* pi = a.pi_lo + t.off_ll;
  const float* gg = a.gg + t.off_ll;
  float* scr = a.scratch + t.off_ll;
  float* u = a.u + (long long)it.x * a.wn;
  const int nblk = (t.DD + kFfTile - 1) >> kFfTileBits;   // low blocks that carry data; the rest of pad(dtheta) is zero
  float x[16];
  if (high || t.ll <= kFfTileBits) {
    for (int k = tid; k < a.wn; k += kFfThreads) bins[k] = 0.f;
  }
  if (!high) {
    const int j0 = (tile << kFfTileBits) + tid;
    const float* dth = a.dtheta + t.off_dd;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int j = j0 + i * 512;
      x[i] = j < t.DD ? dth[j] : 0.f;
    }
    ff_tile_fwht(x, s, 0, t.ll < kFfTileBits ? t.ll : kFfTileBits);
    if (t.ll > kFfTileBits) {
#pragma unroll
      for (int i = 0; i < 16; ++i) scr[j0 + i * 512] = x[i];
      ff_signal(a.counters + it.x);
      return;
    }
    __syncthreads();
    const int LL = 1 << t.ll;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int j = j0 + i * 512;
      if (j < LL) atomicAdd(&bins[pi[j]], x[i] * gg[j]);
    }
    __syncthreads();
    const int n1 = LL < a.wn ? LL : a.wn;
    for (int k = tid; k < n1; k += kFfThreads) atomicAdd(&u[k], bins[k]);
    return;
  }
  const int lc = 2 * kFfTileBits - t.ll;
  const int tc = ff_hi_tconst(tid, lc) + (tile << lc);
  uint16_t pv[16];
  float gv[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {                         // Pi and G do not depend on the first pass: fetch before waiting
    const int j = tc + ff_hi_ioff(i, lc);
    pv[i] = pi[j];
    gv[i] = gg[j];
  }
  ff_wait(a.counters + it.x, nblk);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int j = tc + ff_hi_ioff(i, lc);
    x[i] = (j >> kFfTileBits) < nblk ? __ldcg(scr + j) : 0.f;
  }
  ff_tile_fwht(x, s, lc, kFfTileBits);
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 16; ++i) atomicAdd(&bins[pv[i]], x[i] * gv[i]);
  __syncthreads();
  for (int k = tid; k < a.wn; k += kFfThreads) atomicAdd(&u[k], bins[k]);
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
