// One-shot all-reduce of the flat adapter-gradient bucket over NVLink peer memory, fused with the SGD-momentum update and the
// zeroing of the bucket for the next step.
//
// replaces: dist.all_reduce(flat) + flat.mul_(1/N) + torch.optim.SGD.step() + bucket.zero() of the data-parallel step — the
// reference's optimiser semantics are torch.optim.SGD(momentum, dampening 0, weight_decay, nesterov False) as built by
// few_shot/vision_benchmark/optim/build.py:120-159 and stepped at few_shot/vision_benchmark/evaluation/kadaptation_clip.py:
// 308-314.  Payloads are 0.18 - 13 MB, i.e. latency-bound: every rank reads every peer's bucket directly
// (symmetric-memory peer pointers; `ld.relaxed.sys`, peer lines are never cached in the local L2), sums the ranks in rank
// order — so all ranks compute bit-identical parameters — and updates its own replica.
//
// Cross-GPU synchronisation uses the symmetric-memory signal pads: slot (block, peer) is raised with a system-scope CAS
// 0 -> 1 by the peer and consumed with a CAS 1 -> 0 by the owner, so the protocol re-arms itself and a CUDA-graph replay needs
// no host-side epoch.  Barrier A (before the reads) says "every rank's backward has written its bucket"; barrier B (after the
// reads) says "every rank has read my bucket", after which the bucket may be zeroed.  Each thread block runs its own
// barriers on its own pad slots and owns one contiguous range of the bucket.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace peft {

constexpr int kArMaxWorld = 16;
constexpr int kArThreads = 512;
constexpr int kArMaxBlocks = 32;

struct AllreduceSgdArgs {
  const float* const* peer_grads;   // device array [world]: every rank's flat gradient bucket (own rank included)
  uint32_t* const* peer_pads;       // device array [world]: every rank's signal pad
  float* grad;                      // own bucket (== peer_grads[rank])
  float* param;                     // flat fp32 parameters (views of the nn.Parameters)
  float* momentum;                  // flat momentum buffer (zero-initialised: mu * 0 + g == torch's first-step buf = g)
  const float* lr;                  // device scalar: LR schedules change it without re-capturing the step graph
  long long n;
  int rank, world;
  float mu, weight_decay;
  int zero_grad;                    // 1: leave the bucket zeroed for the next step
  int pad_base;                     // first uint32 slot of the signal pad this kernel may use (2 * blocks * world slots)
};

__device__ __forceinline__ void ar_signal(uint32_t* addr) {
  uint32_t old;
  unsigned spins = 0;
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
    if (old != 0 && ++spins > (1u << 26)) __trap();       // watchdog: a dead peer must not hang this GPU
  } while (old != 0);
}
__device__ __forceinline__ void ar_wait(uint32_t* addr) {
  uint32_t old;
  unsigned spins = 0;
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
    if (old != 1 && ++spins > (1u << 26)) __trap();
  } while (old != 1);
}
// all ranks' block `blockIdx.x` meet; `phase` (0 / 1) selects the slot set so that barrier B cannot consume A's signals
__device__ __forceinline__ void ar_barrier(const AllreduceSgdArgs& a, int phase) {
  __syncthreads();                                       // this block's own work is done ...
  if (threadIdx.x < a.world) {
    const int peer = threadIdx.x;
    const int base = a.pad_base + ((phase * gridDim.x + blockIdx.x) * a.world);
    __threadfence_system();                              // ... and visible system-wide before the peers are told so
    ar_signal(a.peer_pads[peer] + base + a.rank);        // "rank `rank` arrived" in the peer's pad
    ar_wait(a.peer_pads[a.rank] + base + peer);          // the peer arrived (its signal landed in my pad)
  }
  __syncthreads();
}
__device__ __forceinline__ float4 ar_ld_peer(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kArThreads) allreduce_sgd_kernel(const AllreduceSgdArgs a) {
  const long long n4 = a.n >> 2;                          // the bucket is padded to a multiple of 4 floats by the host
  const long long per = (n4 + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = (lo + per < n4) ? lo + per : n4;
  const float lr = *a.lr;
  const float inv = 1.0f / (float)a.world;
  if (a.world > 1) ar_barrier(a, 0);
  for (long long i = lo + threadIdx.x; i < hi; i += kArThreads) {
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.world > 1) {
      for (int r = 0; r < a.world; ++r) {                 // rank order: identical sums on every rank
        const float4 v = ar_ld_peer(a.peer_grads[r] + 4 * i);
        g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
      }
      g.x *= inv; g.y *= inv; g.z *= inv; g.w *= inv;
    } else {
      g = reinterpret_cast<const float4*>(a.grad)[i];
    }
    float4 p = reinterpret_cast<float4*>(a.param)[i];
    float4 m = reinterpret_cast<float4*>(a.momentum)[i];
    // torch.optim.SGD: g += wd * p; buf = mu * buf + g; p -= lr * buf      (dampening 0, nesterov False)
    g.x = fmaf(a.weight_decay, p.x, g.x); g.y = fmaf(a.weight_decay, p.y, g.y);
    g.z = fmaf(a.weight_decay, p.z, g.z); g.w = fmaf(a.weight_decay, p.w, g.w);
    m.x = fmaf(a.mu, m.x, g.x); m.y = fmaf(a.mu, m.y, g.y); m.z = fmaf(a.mu, m.z, g.z); m.w = fmaf(a.mu, m.w, g.w);
    p.x = fmaf(-lr, m.x, p.x); p.y = fmaf(-lr, m.y, p.y); p.z = fmaf(-lr, m.z, p.z); p.w = fmaf(-lr, m.w, p.w);
    reinterpret_cast<float4*>(a.param)[i] = p;
    reinterpret_cast<float4*>(a.momentum)[i] = m;
  }
  if (a.world > 1) ar_barrier(a, 1);                      // every rank has read this block's range of my bucket
  if (a.zero_grad)
    for (long long i = lo + threadIdx.x; i < hi; i += kArThreads) reinterpret_cast<float4*>(a.grad)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

}  // namespace peft
