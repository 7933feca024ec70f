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
//     (read with ld.global.cg).  The forward stores only the intermediate rows below DD and drops their L2 lines after the
//     read (discard.global.L2), so the intermediate never costs a DRAM write-back.
//   * the kernels are persistent (one cooperative launch of occupancy x SMs CTAs, items taken round-robin): while a CTA
//     transforms item k, the Pi / G (or theta0 / dtheta) bytes of its item k + grid are already in flight into a shared-
//     memory staging buffer (cp.async), so the DRAM latency of a tile is hidden behind the butterflies of the previous one.
//     Producers never wait and every item only waits on lower-numbered items, so the spin cannot deadlock while all CTAs are
//     co-resident (guaranteed by the cooperative launch).
// HBM traffic per step is then the algorithmic minimum: 6 B per LL element (uint16 Pi mod wn + fp32 G) plus 8 B (fwd:
// theta0 read, theta write) or 4 B (bwd: dtheta read) per DD element.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/peft_b200.h"
#include "ptx_sm100.cuh"

namespace peft {

constexpr int kFfThreads = 512;
constexpr int kFfTile = 8192;      // elements per tile = 2^13
constexpr int kFfTileBits = 13;
constexpr int kFfMaxWn = 4096;
constexpr int kFfDescCache = 384;  // tensor descriptors kept in shared memory (the rest is read from global)

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
  const int2* items;            // (tensor, tile | kind << 30): kind 0 = low block, 1 = high tile; producers before consumers
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
  int T;
  int n_items;
  int discard;                  // forward: drop the intermediate's L2 lines after the low pass has read them
};

// shared memory of one CTA: tile | w or bins | fp32 staging | uint16 staging | descriptor cache
__host__ __device__ constexpr int ff_smem_bytes(int wn) {
  return (kFfTile + wn + kFfTile) * 4 + kFfTile * 2 + kFfDescCache * (int)sizeof(peft_ff_tensor);
}

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
      if (v < target) __nanosleep(100);
    } while (v < target);
  }
  __syncthreads();
}
__device__ __forceinline__ peft_ff_tensor ff_desc(const FfArgs& a, const peft_ff_tensor* cache, int idx) {
  return idx < kFfDescCache ? cache[idx] : a.tensors[idx];
}
__device__ __forceinline__ void ff_cp16(void* sdst, const void* g, int bytes) {   // bytes < 16: the rest is zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(ptx::smem_u32(sdst)), "l"(g), "r"(bytes) : "memory");
}

// ---- staging (cp.async) ----------------------------------------------------------------------------------------------
// what the NEXT item of this CTA needs from DRAM and can be fetched ahead of its dependencies:
//   mode 1: a high tile's Pi (uint16) and G (fp32), laid out by tile-local element e (needs 2^lc >= 8 columns per row
//           segment and 16-byte aligned tensor starts, i.e. off_ll % 8 == 0);
//   mode 2: a low block's theta0 (forward) / dtheta (backward), zero-filled at and beyond DD (off_dd % 4 == 0);
//   mode 0: nothing staged (single-pass tensors and the odd layouts): the item loads directly.
__device__ __forceinline__ int ff_stage_mode(int2 it, const peft_ff_tensor& t) {
  if (it.y >> 30) return (2 * kFfTileBits - t.ll >= 3 && (t.off_ll & 7) == 0) ? 1 : 0;
  return (t.ll > kFfTileBits && (t.off_dd & 3) == 0) ? 2 : 0;
}
__device__ __forceinline__ void ff_stage(const FfArgs& a, int2 it, const peft_ff_tensor& t, const float* vec, float* sg,
                                         uint16_t* sp) {
  const int tid = threadIdx.x, tile = it.y & 0x3fffffff;
  const int mode = ff_stage_mode(it, t);
  if (mode == 1) {
    const int lc = 2 * kFfTileBits - t.ll;
    const float* gg = a.gg + t.off_ll + (tile << lc);
    const uint16_t* pi = a.pi_lo + t.off_ll + (tile << lc);
#pragma unroll
    for (int q = tid; q < kFfTile / 4; q += kFfThreads) {
      const int e0 = q << 2;
      ff_cp16(sg + e0, gg + ((e0 >> lc) << kFfTileBits) + (e0 & ((1 << lc) - 1)), 16);
    }
#pragma unroll
    for (int q = tid; q < kFfTile / 8; q += kFfThreads) {
      const int e0 = q << 3;
      ff_cp16(sp + e0, pi + ((e0 >> lc) << kFfTileBits) + (e0 & ((1 << lc) - 1)), 16);
    }
  } else if (mode == 2) {
    const float* src = vec + t.off_dd;
#pragma unroll
    for (int q = tid; q < kFfTile / 4; q += kFfThreads) {
      const int j = (tile << kFfTileBits) + (q << 2);
      const int rem = t.DD - j;
      ff_cp16(sg + (q << 2), src + (rem > 0 ? j : 0), rem >= 4 ? 16 : (rem > 0 ? rem * 4 : 0));
    }
  }
  ptx::cp_async_commit();
}

// ---- forward ---------------------------------------------------------------------------------------------------------
// high tile: x[j] = w[Pi[j] mod wn] * G[j] generated on the fly, butterflies over bits [13, ll), fp32 intermediate out
//            (only the rows the low pass will read: row < ceil(DD / 8192))
// low block: butterflies over bits [0, min(ll, 13)), theta = theta0 + coef * x for j < DD
__global__ void __launch_bounds__(kFfThreads, 2) ff_fwd_kernel(FfArgs a) {
  extern __shared__ __align__(16) float ff_smem[];
  float* s = ff_smem;
  float* w = s + kFfTile;
  float* sg = w + a.wn;
  uint16_t* sp = reinterpret_cast<uint16_t*>(sg + kFfTile);
  peft_ff_tensor* cache = reinterpret_cast<peft_ff_tensor*>(sp + kFfTile);
  const int tid = threadIdx.x;
  int k = blockIdx.x;
  if (k >= a.n_items) return;
  {
    const int nc = (a.T < kFfDescCache ? a.T : kFfDescCache) * (int)(sizeof(peft_ff_tensor) / 16);
    for (int i = tid; i < nc; i += kFfThreads) reinterpret_cast<int4*>(cache)[i] = reinterpret_cast<const int4*>(a.tensors)[i];
  }
  __syncthreads();
  int2 it = a.items[k];
  ff_stage(a, it, ff_desc(a, cache, it.x), a.theta0, sg, sp);
  int wcur = -1;                                           // tensor whose w row is in shared memory
  for (;;) {
    const int kn = k + gridDim.x;
    const bool more = kn < a.n_items;
    int2 itn = make_int2(0, 0);
    if (more) itn = a.items[kn];                           // needed only after the staging buffers are free again
    const peft_ff_tensor t = ff_desc(a, cache, it.x);
    const int tile = it.y & 0x3fffffff;
    const bool high = (it.y >> 30) != 0;
    const int mode = ff_stage_mode(it, t);
    const uint16_t* pi = a.pi_lo + t.off_ll;
    const float* gg = a.gg + t.off_ll;
    float* scr = a.scratch + t.off_ll;
    const float* wv = a.wv + (long long)it.x * a.wn;
    const int nblk = (t.DD + kFfTile - 1) >> kFfTileBits;
    float x[16];
    ptx::cp_async_wait<0>();
    __syncthreads();                                       // (A) this item's staged bytes have landed
    if (high) {
      const int lc = 2 * kFfTileBits - t.ll;               // log2(columns per tile)
      const int tc = ff_hi_tconst(tid, lc) + (tile << lc);
      uint16_t pv[16];
      if (mode == 1) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          pv[i] = sp[i * 512 + tid];
          x[i] = sg[i * 512 + tid];
        }
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int j = tc + ff_hi_ioff(i, lc);
          pv[i] = pi[j];
          x[i] = gg[j];
        }
      }
      if (wcur != it.x) {
        for (int q = tid; q < a.wn; q += kFfThreads) w[q] = wv[q];
        wcur = it.x;
      }
      __syncthreads();                                     // (B) staging consumed, w visible
      if (more) ff_stage(a, itn, ff_desc(a, cache, itn.x), a.theta0, sg, sp);
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] *= w[pv[i]];
      ff_tile_fwht(x, s, lc, kFfTileBits);
      if (lc < 9) {                                        // the large tensors: element i sits i * 2^(22-lc) further on
        const int S = 1 << (22 - lc), r0 = tc >> kFfTileBits, rs = S >> kFfTileBits;
        float* p = scr + tc;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (r0 + i * rs < nblk) p[i * S] = x[i];
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int j = tc + ff_hi_ioff(i, lc);
          if ((j >> kFfTileBits) < nblk) scr[j] = x[i];
        }
      }
      ff_signal(a.counters + it.x);
    } else {
      const int j0 = (tile << kFfTileBits) + tid;
      const float* th0 = a.theta0 + t.off_dd;
      float* th = a.theta + t.off_dd;
      float t0v[16];
      if (mode == 2) {
#pragma unroll
        for (int i = 0; i < 16; ++i) t0v[i] = sg[i * 512 + tid];
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int j = j0 + i * 512;
          t0v[i] = j < t.DD ? th0[j] : 0.f;
        }
      }
      __syncthreads();                                     // (B)
      if (more) ff_stage(a, itn, ff_desc(a, cache, itn.x), a.theta0, sg, sp);
      if (t.ll <= kFfTileBits) {                           // single-pass tensor: generate the input here
        const int LL = 1 << t.ll;
        uint16_t pv[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int j = j0 + i * 512;
          pv[i] = j < LL ? pi[j] : (uint16_t)0;
          x[i] = j < LL ? gg[j] : 0.f;
        }
        if (wcur != it.x) {
          for (int q = tid; q < a.wn; q += kFfThreads) w[q] = wv[q];
          wcur = it.x;
        }
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
      if (a.discard && t.ll > kFfTileBits && (tid & 31) == 0 && (t.off_ll & 31) == 0) {
        // every load of this block has been consumed (the transform went through shared memory); each warp read whole
        // 128-byte lines that nobody reads again before the next forward overwrites them
#pragma unroll
        for (int i = 0; i < 16; ++i) asm volatile("discard.global.L2 [%0], 128;" ::"l"(scr + j0 + i * 512) : "memory");
      }
    }
    if (!more) break;
    k = kn;
    it = itn;
  }
}

// ---- backward --------------------------------------------------------------------------------------------------------
// low block: x = pad(dtheta), butterflies over bits [0, min(ll,13)); single-pass tensors bin here, others write fp32
// high tile: butterflies over bits [13, ll), times G, folded through Pi^T into wn bins (shared atomics, then global)
// One CTA per item, 3 CTAs per SM.  A persistent cp.async-staged variant (like the forward) was measured SLOWER here
// (910 vs 850 us on the ViT-B/32 set): the 16 shared-memory float atomics per thread compile to ATOMS.CAST.SPIN loops and
// are the top stall (ncu: 29 % of samples), which a third resident CTA hides better than the staging does.
__global__ void __launch_bounds__(kFfThreads, 3) ff_bwd_kernel(FfArgs a) {
  extern __shared__ float ff_smem[];
  float* s = ff_smem;
  float* bins = ff_smem + kFfTile;
  // Work items are handed out by an atomic TICKET, not by blockIdx: a CTA that spins on a per-tensor counter only ever waits
  // for items with LOWER tickets, and those were taken by CTAs that are already running (producers never wait) — so forward
  // progress does not depend on the order in which the hardware happens to dispatch thread blocks (MPS, preemption, debugger).
  __shared__ int s_ticket;
  if (threadIdx.x == 0) s_ticket = atomicAdd(a.counters + a.T, 1);
  __syncthreads();
  const int2 it = a.items[s_ticket];
  const peft_ff_tensor t = a.tensors[it.x];
  const int tid = threadIdx.x;
  const bool high = (it.y >> 30) != 0;
  // a high item is a RUN of consecutive tiles of one tensor (count - 1 in bits 20..27): they are folded into the same shared-memory
  // bins, so a bin gets one global atomic per run instead of one per tile (the wn bins of a tensor are shared by all its tiles —
  // 512 for an fc1 weight — and same-address atomics serialise in L2)
  const int tile = high ? (it.y & 0xfffff) : (it.y & 0x3fffffff);
  const int ntile = high ? ((it.y >> 20) & 0xff) + 1 : 1;
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
#pragma unroll 1
  for (int q = 0; q < ntile; ++q) {
  const int tc = ff_hi_tconst(tid, lc) + ((tile + q) << lc);
  uint16_t pv[16];
  float gv[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {                         // Pi and G do not depend on the first pass: fetch before waiting
    const int j = tc + ff_hi_ioff(i, lc);
    pv[i] = pi[j];
    gv[i] = gg[j];
  }
  if (q == 0) ff_wait(a.counters + it.x, nblk);
  else __syncthreads();                                  // the previous tile's transform is done with the shared tile
  if (lc < 9) {
    const int S = 1 << (22 - lc), r0 = tc >> kFfTileBits, rs = S >> kFfTileBits;
    const float* p = scr + tc;
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = r0 + i * rs < nblk ? __ldcg(p + i * S) : 0.f;
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int j = tc + ff_hi_ioff(i, lc);
      x[i] = (j >> kFfTileBits) < nblk ? __ldcg(scr + j) : 0.f;
    }
  }
  ff_tile_fwht(x, s, lc, kFfTileBits);
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 16; ++i) atomicAdd(&bins[pv[i]], x[i] * gv[i]);
  }
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
