// Skinny "token-reduction" products for the adapter gradients:
//     Out[I, J] += sum_m X[m, i] * Y[m, j]          I in {16, 32, 64},  J a multiple of 64,  m over all B*L tokens
// e.g. dUp^T = p^T dC, dDn = dp^T x  (LoRA / Kronecker factors), Houlsby dW_up = z^T dy, dW_down = dz^T LN(h).
// These are the only weight-gradient products the frozen-backbone step needs; they are HBM-bound (each token row of
// Y is read once) so the math runs on warp-level mma.sync out of cp.async-staged shared memory and the partial sums of
// the 4 warps of a CTA are combined in shared memory before one fp32 atomicAdd per output element per CTA.
#pragma once
#include "ptx_sm100.cuh"

namespace peft {

constexpr int kRedThreads = 128;
constexpr int kRedRowsPerStage = 64;   // token rows per pipeline stage (16 per warp)
// cp.async ring: three (I = 64: two) stages in flight while one is multiplied — what fits 48 KB of static shared memory.  r02:
// with a two-stage ring and two CTAs per SM only ~24 KB per SM were in flight and the kernel ran at the latency of its loads
// (2.3 TB/s at M = 50 432; 3.6 - 3.8 TB/s with the deeper ring and 4 - 6 CTAs per SM).
template <int IT> struct RedStages { static constexpr int value = IT <= 2 ? 4 : 3; };

__device__ __forceinline__ uint32_t sw_off_x(int row, int chunk, int chunks_per_row) {
  // rows of 32/64/128 bytes: XOR the 16-byte chunk index so 8 consecutive rows hit 8 distinct bank groups
  if (chunks_per_row == 8) return (uint32_t)(row * 128 + ((chunk ^ (row & 7)) << 4));
  if (chunks_per_row == 4) return (uint32_t)(row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4));
  return (uint32_t)(row * 32 + ((chunk ^ ((row >> 2) & 1)) << 4));
}

template <int IT>  // I = 16 * IT
__global__ void __launch_bounds__(kRedThreads)
tn_reduce_kernel(const __nv_bfloat16* __restrict__ X, long long ldx, const __nv_bfloat16* __restrict__ Y, long long ldy,
                 float* __restrict__ out, long long ldout, int M, int rows_per_cta) {
  constexpr int I = 16 * IT;
  constexpr int XCH = I / 8;                       // 16-byte chunks per X row
  constexpr int kXBytes = kRedRowsPerStage * I * 2;
  constexpr int kYBytes = kRedRowsPerStage * 128;
  constexpr int kRedStages = RedStages<IT>::value;
  __shared__ __align__(128) uint8_t sX[kRedStages][kXBytes];
  __shared__ __align__(128) uint8_t sY[kRedStages][kYBytes];
  static_assert(I * 64 * 4 <= kRedStages * kYBytes, "the combine buffer aliases the Y stages");
  float* sRed = reinterpret_cast<float*>(&sY[0][0]);      // used after the main loop only (its last barrier frees the stages)

  pdl_trigger();                     // PDL (common.cuh)
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j0 = blockIdx.x * 64;
  const long long m_begin = (long long)blockIdx.y * rows_per_cta;
  long long m_end = m_begin + rows_per_cta;
  if (m_end > M) m_end = M;
  const int nstages = (int)((m_end - m_begin + kRedRowsPerStage - 1) / kRedRowsPerStage);

  auto issue = [&](int st, int buf) {       // always commits a group (an empty one past the end) so the wait count is uniform
    if (st >= nstages) { ptx::cp_async_commit(); return; }
    const long long r0 = m_begin + (long long)st * kRedRowsPerStage;
    for (int idx = threadIdx.x; idx < kRedRowsPerStage * XCH; idx += kRedThreads) {
      const int r = idx / XCH, c = idx % XCH;
      const bool ok = (r0 + r) < m_end;
      ptx::cp_async_16(&sX[buf][sw_off_x(r, c, XCH)], X + (ok ? (r0 + r) : 0) * ldx + c * 8, ok);
    }
    for (int idx = threadIdx.x; idx < kRedRowsPerStage * 8; idx += kRedThreads) {
      const int r = idx >> 3, c = idx & 7;
      const bool ok = (r0 + r) < m_end;
      ptx::cp_async_16(&sY[buf][sw_off_x(r, c, 8)], Y + (ok ? (r0 + r) : 0) * ldy + j0 + c * 8, ok);
    }
    ptx::cp_async_commit();
  };

  float acc[IT][8][4];
#pragma unroll
  for (int a = 0; a < IT; ++a)
#pragma unroll
    for (int n = 0; n < 8; ++n) acc[a][n][0] = acc[a][n][1] = acc[a][n][2] = acc[a][n][3] = 0.f;

#pragma unroll
  for (int st = 0; st < kRedStages - 1; ++st) issue(st, st);
  for (int st = 0; st < nstages; ++st) {
    const int buf = st % kRedStages;
    issue(st + kRedStages - 1, (st + kRedStages - 1) % kRedStages);   // its buffer was released by the barrier that ended stage st - 1
    ptx::cp_async_wait<kRedStages - 1>();                             // all but the newest kRedStages - 1 groups: stage st has landed
    __syncthreads();
    const uint32_t uX = ptx::smem_u32(sX[buf]), uY = ptx::smem_u32(sY[buf]);
    const int k0 = warp * 16;  // this warp's 16 token rows of the stage
    const int mat = lane >> 3, r = lane & 7;
    uint32_t bfr[4][4];
#pragma unroll
    for (int nt = 0; nt < 8; nt += 2)
      ptx::ldmatrix_x4_trans(bfr[nt >> 1], uY + sw_off_x(k0 + (mat & 1) * 8 + r, nt + (mat >> 1), 8));
#pragma unroll
    for (int a = 0; a < IT; ++a) {
      uint32_t af[4];
      // A[i][m] = X[m][i]: transposed 8x8 loads; matrices (i lo, m lo), (i hi, m lo), (i lo, m hi), (i hi, m hi)
      ptx::ldmatrix_x4_trans(af, uX + sw_off_x(k0 + (mat >> 1) * 8 + r, a * 2 + (mat & 1), XCH));
#pragma unroll
      for (int nt = 0; nt < 8; nt += 2) {
        const uint32_t b0[2] = {bfr[nt >> 1][0], bfr[nt >> 1][1]}, b1[2] = {bfr[nt >> 1][2], bfr[nt >> 1][3]};
        ptx::mma_bf16_16816(acc[a][nt], af, b0);
        ptx::mma_bf16_16816(acc[a][nt + 1], af, b1);
      }
    }
    __syncthreads();
  }

  // combine the 4 warps' partial sums in shared memory — warp after warp with plain read-modify-writes (a float atomicAdd
  // on shared memory is a CAS loop, and the four warps collide on every address) — then one global atomic per element
#pragma unroll 1
  for (int w = 0; w < kRedThreads / 32; ++w) {
    if (warp == w) {
#pragma unroll
      for (int a = 0; a < IT; ++a)
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const int row = a * 16 + (lane >> 2), col = nt * 8 + (lane & 3) * 2;
          float2* lo = reinterpret_cast<float2*>(&sRed[row * 64 + col]);
          float2* hi = reinterpret_cast<float2*>(&sRed[(row + 8) * 64 + col]);
          if (w == 0) {
            *lo = make_float2(acc[a][nt][0], acc[a][nt][1]);
            *hi = make_float2(acc[a][nt][2], acc[a][nt][3]);
          } else {
            const float2 l = *lo, h = *hi;
            *lo = make_float2(l.x + acc[a][nt][0], l.y + acc[a][nt][1]);
            *hi = make_float2(h.x + acc[a][nt][2], h.y + acc[a][nt][3]);
          }
        }
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < I * 64; i += kRedThreads) {
    const int row = i >> 6, col = i & 63;
    atomicAdd(out + (long long)row * ldout + j0 + col, sRed[i]);
  }
}

}  // namespace peft
