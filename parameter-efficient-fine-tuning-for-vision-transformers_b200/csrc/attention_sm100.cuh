// Fused flash-style multi-head attention for the ViT sequence lengths (L = 50 / 197 / 257, head_dim = 64), forward and
// backward, reading q/k/v straight out of the packed [B*L, 3D] projection buffer and writing [B*L, D] / [B*L, 3D]:
// the head split / permute / contiguous copies of the reference and the (B*H, L, L) score tensors
// (R6: FS/kadaptation.py:223-235, FS/model.py:990-999) never touch HBM.
//
// v1 uses warp-level mma.sync (m16n8k16, bf16 in / fp32 accumulate) with the whole K/V (or Q/dO) of one head resident in
// shared memory (<= 320 rows), 64-row tiles per CTA, online softmax in exp2 domain.
//   fwd      : grid (ceil(L/64), B*H)  -> O, lse2 (= rowmax*c + log2(rowsum), c = scale*log2 e)
//   bwd dQ   : grid (ceil(L/64), B*H)  -> dQ, delta (= rowsum(dO*O))
//   bwd dKdV : grid (ceil(L/64), B*H)  -> dK, dV           (recomputes S^T; no atomics, deterministic)
#pragma once
#include "ptx_sm100.cuh"

namespace peft {

constexpr int kAttnThreads = 128;
constexpr int kHd = 64;

// 64-column bf16 rows (128 B) with a 16-byte-chunk XOR swizzle: conflict-free ldmatrix.
__device__ __forceinline__ uint32_t sw_off(int row, int chunk) { return (uint32_t)(row * 128 + ((chunk ^ (row & 7)) << 4)); }

// Load `rows` rows x 64 bf16 from global (row pitch ld elements) into swizzled smem; rows >= valid are zero-filled.
__device__ __forceinline__ void load_rows_async(uint8_t* smem, const __nv_bfloat16* g, long long ld, int rows, int valid) {
  for (int idx = threadIdx.x; idx < rows * 8; idx += kAttnThreads) {
    const int r = idx >> 3, c = idx & 7;
    const bool ok = r < valid;
    ptx::cp_async_16(smem + sw_off(r, c), g + (long long)(ok ? r : 0) * ld + c * 8, ok);
  }
}

// A-operand fragments (16 rows x 64 k) of a row-major smem tile, rows [r0, r0+16): 4 k-steps x 4 regs.
__device__ __forceinline__ void load_a_frags(uint32_t (&a)[4][4], uint32_t sbase, int r0, int lane) {
  const int m = lane >> 3, r = lane & 7;
  const int row = r0 + (m & 1) * 8 + r;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) ptx::ldmatrix_x4(a[ks], sbase + sw_off(row, ks * 2 + (m >> 1)));
}
// B fragments for C = A * T^T where T is [n][k] row-major in smem (k contiguous): two n-tiles (16 n) x one k16.
__device__ __forceinline__ void load_b_nk(uint32_t (&b)[4], uint32_t sbase, int n0, int kchunk, int lane) {
  const int m = lane >> 3, r = lane & 7;
  ptx::ldmatrix_x4(b, sbase + sw_off(n0 + (m >> 1) * 8 + r, kchunk + (m & 1)));
}
// B fragments for C = A * T where T is [k][n] row-major in smem (n contiguous): two n-tiles (16 n) x one k16.
__device__ __forceinline__ void load_b_kn(uint32_t (&b)[4], uint32_t sbase, int k0, int nchunk, int lane) {
  const int m = lane >> 3, r = lane & 7;
  ptx::ldmatrix_x4_trans(b, sbase + sw_off(k0 + (m & 1) * 8 + r, nchunk + (m >> 1)));
}

struct AttnParams {
  const __nv_bfloat16* qkv;  // [B*L, ldqkv]; q at col h*64, k at D + h*64, v at 2D + h*64
  long long ldqkv;
  __nv_bfloat16* o;          // [B*L, ldo] (fwd out / bwd in)
  long long ldo;
  const __nv_bfloat16* d_o;  // [B*L, lddo]
  long long lddo;
  __nv_bfloat16* dqkv;       // [B*L, lddqkv]
  long long lddqkv;
  float* lse2;               // [B*H, L]
  float* delta;              // [B*H, L]
  int B, H, L, D;
  float scale;
};

// ------------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(kAttnThreads)
attn_fwd_kernel(const AttnParams p, int Lp) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sQ = smem;                   // 64 rows
  uint8_t* sK = sQ + 64 * 128;          // Lp rows
  uint8_t* sV = sK + Lp * 128;          // Lp rows
  const int bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
  const int q0 = blockIdx.x * 64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const __nv_bfloat16* base = p.qkv + (long long)b * p.L * p.ldqkv + h * kHd;
  load_rows_async(sQ, base + (long long)q0 * p.ldqkv, p.ldqkv, 64, p.L - q0);
  load_rows_async(sK, base + p.D, p.ldqkv, Lp, p.L);
  load_rows_async(sV, base + 2 * p.D, p.ldqkv, Lp, p.L);
  ptx::cp_async_commit();
  ptx::cp_async_wait<0>();
  __syncthreads();

  const uint32_t uQ = ptx::smem_u32(sQ), uK = ptx::smem_u32(sK), uV = ptx::smem_u32(sV);
  uint32_t qa[4][4];
  load_a_frags(qa, uQ, warp * 16, lane);
  const float c = p.scale * 1.4426950408889634f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  float oacc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) oacc[i][0] = oacc[i][1] = oacc[i][2] = oacc[i][3] = 0.f;

  for (int kb = 0; kb < Lp; kb += 64) {
    float s[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int nt = 0; nt < 8; nt += 2) {
        uint32_t bf[4];
        load_b_nk(bf, uK, kb + nt * 8, ks * 2, lane);
        const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
        ptx::mma_bf16_16816(s[nt], qa[ks], b0);
        ptx::mma_bf16_16816(s[nt + 1], qa[ks], b1);
      }
    }
    float mx0 = m0, mx1 = m1;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int col = kb + nt * 8 + (lane & 3) * 2;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool ok = (col + (j & 1)) < p.L;
        s[nt][j] = ok ? s[nt][j] * c : -INFINITY;
      }
      mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
      mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float a0 = exp2f(m0 - mx0), a1 = exp2f(m1 - mx1);  // first block: exp2(-inf) = 0
    m0 = mx0; m1 = mx1;
    float r0 = 0.f, r1 = 0.f;
    uint32_t pa[4][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = exp2f(s[nt][0] - m0); s[nt][1] = exp2f(s[nt][1] - m0);
      s[nt][2] = exp2f(s[nt][2] - m1); s[nt][3] = exp2f(s[nt][3] - m1);
      r0 += s[nt][0] + s[nt][1];
      r1 += s[nt][2] + s[nt][3];
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      pa[t][0] = ptx::pack_bf16x2(s[2 * t][0], s[2 * t][1]);
      pa[t][1] = ptx::pack_bf16x2(s[2 * t][2], s[2 * t][3]);
      pa[t][2] = ptx::pack_bf16x2(s[2 * t + 1][0], s[2 * t + 1][1]);
      pa[t][3] = ptx::pack_bf16x2(s[2 * t + 1][2], s[2 * t + 1][3]);
    }
    l0 = l0 * a0 + r0;
    l1 = l1 * a1 + r1;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      oacc[nt][0] *= a0; oacc[nt][1] *= a0; oacc[nt][2] *= a1; oacc[nt][3] *= a1;
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {       // k16 step over keys kb + 16 t
#pragma unroll
      for (int nt = 0; nt < 8; nt += 2) {
        uint32_t bf[4];
        load_b_kn(bf, uV, kb + t * 16, nt, lane);
        const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
        ptx::mma_bf16_16816(oacc[nt], pa[t], b0);
        ptx::mma_bf16_16816(oacc[nt + 1], pa[t], b1);
      }
    }
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = 1.f / l0, i1 = 1.f / l1;
  const int row0 = q0 + warp * 16 + (lane >> 2), row1 = row0 + 8;
  __nv_bfloat16* ob = p.o + (long long)b * p.L * p.ldo + h * kHd;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    const int col = nt * 8 + (lane & 3) * 2;
    if (row0 < p.L) *reinterpret_cast<uint32_t*>(ob + (long long)row0 * p.ldo + col) = ptx::pack_bf16x2(oacc[nt][0] * i0, oacc[nt][1] * i0);
    if (row1 < p.L) *reinterpret_cast<uint32_t*>(ob + (long long)row1 * p.ldo + col) = ptx::pack_bf16x2(oacc[nt][2] * i1, oacc[nt][3] * i1);
  }
  if ((lane & 3) == 0) {
    if (row0 < p.L) p.lse2[(long long)bh * p.L + row0] = m0 + log2f(l0);
    if (row1 < p.L) p.lse2[(long long)bh * p.L + row1] = m1 + log2f(l1);
  }
}

// ------------------------------------------------------------------------------------------------ backward: dQ (+delta)
__global__ void __launch_bounds__(kAttnThreads)
attn_bwd_dq_kernel(const AttnParams p, int Lp) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sQ = smem;                   // 64 rows
  uint8_t* sdO = sQ + 64 * 128;         // 64 rows
  uint8_t* sK = sdO + 64 * 128;         // Lp rows
  uint8_t* sV = sK + Lp * 128;          // Lp rows
  const int bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
  const int q0 = blockIdx.x * 64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const __nv_bfloat16* base = p.qkv + (long long)b * p.L * p.ldqkv + h * kHd;
  load_rows_async(sQ, base + (long long)q0 * p.ldqkv, p.ldqkv, 64, p.L - q0);
  load_rows_async(sdO, p.d_o + ((long long)b * p.L + q0) * p.lddo + h * kHd, p.lddo, 64, p.L - q0);
  load_rows_async(sK, base + p.D, p.ldqkv, Lp, p.L);
  load_rows_async(sV, base + 2 * p.D, p.ldqkv, Lp, p.L);
  ptx::cp_async_commit();

  // delta = rowsum(dO * O) for this warp's two rows per thread, straight from global (overlaps the async copies)
  const int row0 = q0 + warp * 16 + (lane >> 2), row1 = row0 + 8;
  float d0 = 0.f, d1 = 0.f;
  {
    const __nv_bfloat16* ob = p.o + (long long)b * p.L * p.ldo + h * kHd;
    const __nv_bfloat16* gb = p.d_o + (long long)b * p.L * p.lddo + h * kHd;
    const int cq = (lane & 3) * 16;  // each of the 4 lanes of a quad sums 16 columns
    if (row0 < p.L) {
#pragma unroll
      for (int j = 0; j < 16; j += 8) {
        const uint4 a = *reinterpret_cast<const uint4*>(ob + (long long)row0 * p.ldo + cq + j);
        const uint4 g = *reinterpret_cast<const uint4*>(gb + (long long)row0 * p.lddo + cq + j);
        const uint32_t* au = &a.x; const uint32_t* gu = &g.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) { const float2 x = ptx::unpack_bf16x2(au[k]), y = ptx::unpack_bf16x2(gu[k]); d0 += x.x * y.x + x.y * y.y; }
      }
    }
    if (row1 < p.L) {
#pragma unroll
      for (int j = 0; j < 16; j += 8) {
        const uint4 a = *reinterpret_cast<const uint4*>(ob + (long long)row1 * p.ldo + cq + j);
        const uint4 g = *reinterpret_cast<const uint4*>(gb + (long long)row1 * p.lddo + cq + j);
        const uint32_t* au = &a.x; const uint32_t* gu = &g.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) { const float2 x = ptx::unpack_bf16x2(au[k]), y = ptx::unpack_bf16x2(gu[k]); d1 += x.x * y.x + x.y * y.y; }
      }
    }
    d0 += __shfl_xor_sync(0xffffffffu, d0, 1); d0 += __shfl_xor_sync(0xffffffffu, d0, 2);
    d1 += __shfl_xor_sync(0xffffffffu, d1, 1); d1 += __shfl_xor_sync(0xffffffffu, d1, 2);
    if ((lane & 3) == 0) {
      if (row0 < p.L) p.delta[(long long)bh * p.L + row0] = d0;
      if (row1 < p.L) p.delta[(long long)bh * p.L + row1] = d1;
    }
  }
  const float ls0 = row0 < p.L ? p.lse2[(long long)bh * p.L + row0] : 0.f;
  const float ls1 = row1 < p.L ? p.lse2[(long long)bh * p.L + row1] : 0.f;
  ptx::cp_async_wait<0>();
  __syncthreads();

  const uint32_t uQ = ptx::smem_u32(sQ), udO = ptx::smem_u32(sdO), uK = ptx::smem_u32(sK), uV = ptx::smem_u32(sV);
  uint32_t qa[4][4], ga[4][4];
  load_a_frags(qa, uQ, warp * 16, lane);
  load_a_frags(ga, udO, warp * 16, lane);
  const float c = p.scale * 1.4426950408889634f;
  float dq[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;

  for (int kb = 0; kb < Lp; kb += 64) {
    float s[8][4], dp[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; dp[i][0] = dp[i][1] = dp[i][2] = dp[i][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int nt = 0; nt < 8; nt += 2) {
        uint32_t bf[4];
        load_b_nk(bf, uK, kb + nt * 8, ks * 2, lane);
        { const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
          ptx::mma_bf16_16816(s[nt], qa[ks], b0); ptx::mma_bf16_16816(s[nt + 1], qa[ks], b1); }
        load_b_nk(bf, uV, kb + nt * 8, ks * 2, lane);
        { const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
          ptx::mma_bf16_16816(dp[nt], ga[ks], b0); ptx::mma_bf16_16816(dp[nt + 1], ga[ks], b1); }
      }
    }
    uint32_t dsa[4][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int col = kb + nt * 8 + (lane & 3) * 2;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool ok = (col + (j & 1)) < p.L;
        const float pr = ok ? exp2f(s[nt][j] * c - ((j < 2) ? ls0 : ls1)) : 0.f;
        s[nt][j] = pr * (dp[nt][j] - ((j < 2) ? d0 : d1)) * p.scale;   // dS (already times scale)
      }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      dsa[t][0] = ptx::pack_bf16x2(s[2 * t][0], s[2 * t][1]);
      dsa[t][1] = ptx::pack_bf16x2(s[2 * t][2], s[2 * t][3]);
      dsa[t][2] = ptx::pack_bf16x2(s[2 * t + 1][0], s[2 * t + 1][1]);
      dsa[t][3] = ptx::pack_bf16x2(s[2 * t + 1][2], s[2 * t + 1][3]);
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
#pragma unroll
      for (int nt = 0; nt < 8; nt += 2) {
        uint32_t bf[4];
        load_b_kn(bf, uK, kb + t * 16, nt, lane);
        const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
        ptx::mma_bf16_16816(dq[nt], dsa[t], b0);
        ptx::mma_bf16_16816(dq[nt + 1], dsa[t], b1);
      }
    }
  }
  __nv_bfloat16* ob = p.dqkv + (long long)b * p.L * p.lddqkv + h * kHd;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    const int col = nt * 8 + (lane & 3) * 2;
    if (row0 < p.L) *reinterpret_cast<uint32_t*>(ob + (long long)row0 * p.lddqkv + col) = ptx::pack_bf16x2(dq[nt][0], dq[nt][1]);
    if (row1 < p.L) *reinterpret_cast<uint32_t*>(ob + (long long)row1 * p.lddqkv + col) = ptx::pack_bf16x2(dq[nt][2], dq[nt][3]);
  }
}

// ------------------------------------------------------------------------------------------------ backward: dK, dV
// The CTA owns 64 keys (16 per warp) and sweeps all query blocks with the transposed products
//   S^T = K Q^T,  P^T = exp2(S^T c - lse2[q]),  dV += P^T dO,  dP^T = V dO^T,  dS^T = P^T (dP^T - delta[q]) scale,
//   dK += dS^T Q.
__global__ void __launch_bounds__(kAttnThreads)
attn_bwd_dkv_kernel(const AttnParams p, int Lp) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sK = smem;                   // 64 rows (own keys)
  uint8_t* sV = sK + 64 * 128;          // 64 rows
  uint8_t* sQ = sV + 64 * 128;          // Lp rows
  uint8_t* sdO = sQ + Lp * 128;         // Lp rows
  float* sLse = reinterpret_cast<float*>(sdO + Lp * 128);  // Lp
  float* sDelta = sLse + Lp;                                // Lp
  const int bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
  const int k0 = blockIdx.x * 64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const __nv_bfloat16* base = p.qkv + (long long)b * p.L * p.ldqkv + h * kHd;
  load_rows_async(sK, base + p.D + (long long)k0 * p.ldqkv, p.ldqkv, 64, p.L - k0);
  load_rows_async(sV, base + 2 * p.D + (long long)k0 * p.ldqkv, p.ldqkv, 64, p.L - k0);
  load_rows_async(sQ, base, p.ldqkv, Lp, p.L);
  load_rows_async(sdO, p.d_o + (long long)b * p.L * p.lddo + h * kHd, p.lddo, Lp, p.L);
  ptx::cp_async_commit();
  for (int i = threadIdx.x; i < Lp; i += kAttnThreads) {
    sLse[i] = i < p.L ? p.lse2[(long long)bh * p.L + i] : 0.f;
    sDelta[i] = i < p.L ? p.delta[(long long)bh * p.L + i] : 0.f;
  }
  ptx::cp_async_wait<0>();
  __syncthreads();

  const uint32_t uK = ptx::smem_u32(sK), uV = ptx::smem_u32(sV), uQ = ptx::smem_u32(sQ), udO = ptx::smem_u32(sdO);
  uint32_t ka[4][4], va[4][4];
  load_a_frags(ka, uK, warp * 16, lane);
  load_a_frags(va, uV, warp * 16, lane);
  const float c = p.scale * 1.4426950408889634f;
  const int key0 = k0 + warp * 16 + (lane >> 2), key1 = key0 + 8;
  float dk[8][4], dv[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) { dk[i][0] = dk[i][1] = dk[i][2] = dk[i][3] = 0.f; dv[i][0] = dv[i][1] = dv[i][2] = dv[i][3] = 0.f; }

  for (int qb = 0; qb < Lp; qb += 64) {
    float s[8][4], dp[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; dp[i][0] = dp[i][1] = dp[i][2] = dp[i][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int nt = 0; nt < 8; nt += 2) {
        uint32_t bf[4];
        load_b_nk(bf, uQ, qb + nt * 8, ks * 2, lane);
        { const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
          ptx::mma_bf16_16816(s[nt], ka[ks], b0); ptx::mma_bf16_16816(s[nt + 1], ka[ks], b1); }
        load_b_nk(bf, udO, qb + nt * 8, ks * 2, lane);
        { const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
          ptx::mma_bf16_16816(dp[nt], va[ks], b0); ptx::mma_bf16_16816(dp[nt + 1], va[ks], b1); }
      }
    }
    uint32_t pa[4][4], dsa[4][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int qc = qb + nt * 8 + (lane & 3) * 2;
      const float2 ls = *reinterpret_cast<const float2*>(sLse + qc);
      const float2 dl = *reinterpret_cast<const float2*>(sDelta + qc);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int q = qc + (j & 1);
        const int key = (j < 2) ? key0 : key1;
        const bool ok = (q < p.L) && (key < p.L);
        const float pr = ok ? exp2f(s[nt][j] * c - ((j & 1) ? ls.y : ls.x)) : 0.f;
        s[nt][j] = pr;
        dp[nt][j] = pr * (dp[nt][j] - ((j & 1) ? dl.y : dl.x)) * p.scale;
      }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      pa[t][0] = ptx::pack_bf16x2(s[2 * t][0], s[2 * t][1]);
      pa[t][1] = ptx::pack_bf16x2(s[2 * t][2], s[2 * t][3]);
      pa[t][2] = ptx::pack_bf16x2(s[2 * t + 1][0], s[2 * t + 1][1]);
      pa[t][3] = ptx::pack_bf16x2(s[2 * t + 1][2], s[2 * t + 1][3]);
      dsa[t][0] = ptx::pack_bf16x2(dp[2 * t][0], dp[2 * t][1]);
      dsa[t][1] = ptx::pack_bf16x2(dp[2 * t][2], dp[2 * t][3]);
      dsa[t][2] = ptx::pack_bf16x2(dp[2 * t + 1][0], dp[2 * t + 1][1]);
      dsa[t][3] = ptx::pack_bf16x2(dp[2 * t + 1][2], dp[2 * t + 1][3]);
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {       // k16 over queries qb + 16 t
#pragma unroll
      for (int nt = 0; nt < 8; nt += 2) {
        uint32_t bf[4];
        load_b_kn(bf, udO, qb + t * 16, nt, lane);
        { const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
          ptx::mma_bf16_16816(dv[nt], pa[t], b0); ptx::mma_bf16_16816(dv[nt + 1], pa[t], b1); }
        load_b_kn(bf, uQ, qb + t * 16, nt, lane);
        { const uint32_t b0[2] = {bf[0], bf[1]}, b1[2] = {bf[2], bf[3]};
          ptx::mma_bf16_16816(dk[nt], dsa[t], b0); ptx::mma_bf16_16816(dk[nt + 1], dsa[t], b1); }
      }
    }
  }
  __nv_bfloat16* okb = p.dqkv + (long long)b * p.L * p.lddqkv + p.D + h * kHd;
  __nv_bfloat16* ovb = okb + p.D;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    const int col = nt * 8 + (lane & 3) * 2;
    if (key0 < p.L) {
      *reinterpret_cast<uint32_t*>(okb + (long long)key0 * p.lddqkv + col) = ptx::pack_bf16x2(dk[nt][0], dk[nt][1]);
      *reinterpret_cast<uint32_t*>(ovb + (long long)key0 * p.lddqkv + col) = ptx::pack_bf16x2(dv[nt][0], dv[nt][1]);
    }
    if (key1 < p.L) {
      *reinterpret_cast<uint32_t*>(okb + (long long)key1 * p.lddqkv + col) = ptx::pack_bf16x2(dk[nt][2], dk[nt][3]);
      *reinterpret_cast<uint32_t*>(ovb + (long long)key1 * p.lddqkv + col) = ptx::pack_bf16x2(dv[nt][2], dv[nt][3]);
    }
  }
}

}  // namespace peft
