// Reshape-trick operands of the small-n KAdaptation update for ALL layers in one launch per direction.
//
// replaces: the chain of ~40 tiny ATen launches (stack / einsum / eye / zeros / cat and their autograd) that built
//     Dn [dep, 2 n r n, n K]   Dn[l, (s,i,r,c), (a,k)]        = A_i[a,c] u^s_i[k,r]            (timm_vit.lowrank_factors_kron,
//     UpT[dep, 2 n r n, 3 n P] UpT[l, (s,i,r,c), (slot_s,c,p)] = v^s_i[r,p], zero elsewhere     FS/kadaptation.py:81-115)
// from the reference-named parameters (phm_rule A [dep,n,n,n], W_left{1,2} u^s [dep,n,K,r], W_right{1,2} v^s [dep,n,r,P]; s = 0: q
// projection -> column slot 0, s = 1: v projection -> column slot 2), and its transpose
//     du^s_i[k,r] = sum_{a,c} A_i[a,c] dDn[l,(s,i,r,c),(a,k)]        dv^s_i[r,p] = sum_c dUpT[l,(s,i,r,c),(slot_s,c,p)]
// and, when phm_rule is trained too, dA_i[a,c] = sum_{s,r,k} u^s_i[k,r] dDn[l,(s,i,r,c),(a,k)] (kron_factors_drule_kernel).
// A step of the fixed-global-batch split is ~4.6 ms, so 40 launches of 2.5 us each were 2 % of it.
#pragma once
#include <cuda_runtime.h>

namespace peft {

__global__ void __launch_bounds__(256) kron_factors_fwd_kernel(const float* __restrict__ rule, const float* __restrict__ wl1,
                                                               const float* __restrict__ wr1, const float* __restrict__ wl2,
                                                               const float* __restrict__ wr2, float* __restrict__ Dn,
                                                               float* __restrict__ UpT, int n, int K, int r, int P) {
  const int nrn = n * r * n;
  const int R = blockIdx.x;                 // row (s, i, rr, c) of both outputs
  const int l = blockIdx.y;
  const int s = R / nrn, rem = R % nrn;
  const int i = rem / (r * n), rr = (rem / n) % r, c = rem % n;
  const float* wl = (s == 0 ? wl1 : wl2) + ((long long)l * n + i) * K * r;      // [K, r]
  const float* wr = (s == 0 ? wr1 : wr2) + (((long long)l * n + i) * r + rr) * P;   // [P]
  const float* A = rule + ((long long)l * n + i) * n * n;                        // [a, c]
  const int D = n * K, D3 = 3 * n * P;
  float* dn = Dn + ((long long)l * 2 * nrn + R) * D;
  float* up = UpT + ((long long)l * 2 * nrn + R) * D3;
  for (int col = threadIdx.x; col < D; col += blockDim.x) {
    const int a = col / K, k = col % K;
    dn[col] = A[a * n + c] * wl[k * r + rr];
  }
  const int lo = (s == 0 ? 0 : 2 * n * P) + c * P;      // the only non-zero run of this row
  for (int col = threadIdx.x; col < D3; col += blockDim.x) up[col] = (col >= lo && col < lo + P) ? wr[col - lo] : 0.f;
}

// one thread per gradient element: [s][l][i][k][rr] of dW_left (n*n terms), then [s][l][i][rr][p] of dW_right (n terms)
__global__ void __launch_bounds__(256) kron_factors_bwd_kernel(const float* __restrict__ rule, const float* __restrict__ dDn,
                                                               const float* __restrict__ dUpT, float* __restrict__ dwl1,
                                                               float* __restrict__ dwr1, float* __restrict__ dwl2,
                                                               float* __restrict__ dwr2, int dep, int n, int K, int r, int P) {
  const int nrn = n * r * n, D = n * K, D3 = 3 * n * P;
  const long long nl = (long long)dep * n * K * r, nr = (long long)dep * n * r * P;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < 2 * nl) {
    const int s = (int)(t / nl);
    long long e = t % nl;
    const int rr = (int)(e % r); e /= r;
    const int k = (int)(e % K); e /= K;
    const int i = (int)(e % n);
    const int l = (int)(e / n);
    const float* A = rule + ((long long)l * n + i) * n * n;
    const float* g = dDn + ((long long)l * 2 * nrn + (long long)s * nrn + (i * r + rr) * n) * D + k;
    float acc = 0.f;
    for (int c = 0; c < n; ++c)
      for (int a = 0; a < n; ++a) acc = fmaf(A[a * n + c], g[(long long)c * D + a * K], acc);
    (s == 0 ? dwl1 : dwl2)[t % nl] = acc;
  } else if (t < 2 * nl + 2 * nr) {
    const long long q = t - 2 * nl;
    const int s = (int)(q / nr);
    long long e = q % nr;
    const int p = (int)(e % P); e /= P;
    const int rr = (int)(e % r); e /= r;
    const int i = (int)(e % n);
    const int l = (int)(e / n);
    const float* g = dUpT + ((long long)l * 2 * nrn + (long long)s * nrn + (i * r + rr) * n) * D3 + (s == 0 ? 0 : 2 * n * P) + p;
    float acc = 0.f;
    for (int c = 0; c < n; ++c) acc += g[(long long)c * D3 + c * P];
    (s == 0 ? dwr1 : dwr2)[q % nr] = acc;
  }
}

// gradient of a TRAINABLE phm_rule: dA_i[a,c] = sum_{s,rr,k} u^s_i[k,rr] dDn[l,(s,i,rr,c),(a,k)]; one warp per (l, i, a, c)
__global__ void __launch_bounds__(256) kron_factors_drule_kernel(const float* __restrict__ dDn, const float* __restrict__ wl1,
                                                                 const float* __restrict__ wl2, float* __restrict__ drule, int dep,
                                                                 int n, int K, int r) {
  const int nrn = n * r * n, D = n * K;
  const int w = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (w >= dep * n * n * n) return;
  const int c = w % n, a = (w / n) % n, i = (w / (n * n)) % n, l = w / (n * n * n);
  float acc = 0.f;
  for (int s = 0; s < 2; ++s) {
    const float* wl = (s == 0 ? wl1 : wl2) + ((long long)l * n + i) * K * r;
    for (int rr = 0; rr < r; ++rr) {
      const float* g = dDn + ((long long)l * 2 * nrn + (long long)s * nrn + (i * r + rr) * n + c) * D + a * K;
      for (int k = lane; k < K; k += 32) acc = fmaf(wl[k * r + rr], g[k], acc);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) drule[w] = acc;
}

}  // namespace peft
