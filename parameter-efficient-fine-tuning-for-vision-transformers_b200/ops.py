"""Tensor-level wrappers over the C ABI (include/peft_b200.h).  torch is used only for device memory and streams.

Every function enqueues on the current torch CUDA stream and returns immediately.  Inputs must be CUDA tensors whose
last dimension is contiguous; 2-D views with a row pitch (column slices of wider buffers) are accepted and passed as
(pointer, ld).  Nothing here falls back to torch math.
"""
from __future__ import annotations

import torch

from . import _lib

EPI_STORE, EPI_GELU, EPI_QUICKGELU, EPI_RESID_F32, EPI_DGELU, EPI_DQUICKGELU, EPI_STORE_F32, EPI_RELU, EPI_DRELU, \
    EPI_GELU_TANH, EPI_DGELU_TANH, EPI_RESID_BF16 = range(12)


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return 0 if t is None else t.data_ptr()


def _mat(t: torch.Tensor, dtype=None, name="tensor"):
    if not t.is_cuda:
        raise RuntimeError(f"peft_b200: {name} must be a CUDA tensor (no CPU fallback)")
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"peft_b200: {name} must be 2-D with a contiguous last dim, got shape {tuple(t.shape)} strides {t.stride()}")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"peft_b200: {name} must be {dtype}, got {t.dtype}")
    return t.data_ptr(), t.stride(0)


def _rowscale(rs, M):
    if rs is None:
        return 0
    if not rs.is_cuda or rs.dtype != torch.float32 or rs.numel() != M or not rs.is_contiguous():
        raise ValueError("gemm: rowscale must be a contiguous CUDA fp32 vector of length M")
    return rs.data_ptr()


def gemm(a, b, out, bias=None, epilogue=EPI_STORE, out2=None, aux=None, rowscale=None):
    """out[M,N] = epilogue((a[M,K] @ b[N,K]^T (+ bias)) * rowscale[:, None]); bf16 operands, see peft_gemm_bf16 / _rs.  fp32
    operands select the fp32 verification kernels (peft_gemm_f32: every tensor fp32, same epilogue semantics)."""
    if a.dtype == torch.float32:
        return _gemm_f32(a, b, out, bias, epilogue, out2, aux, rowscale)
    pa, lda = _mat(a, torch.bfloat16, "a")
    pb, ldb = _mat(b, torch.bfloat16, "b")
    M, K = a.shape
    N = b.shape[0]
    if b.shape[1] != K:
        raise ValueError(f"gemm: K mismatch {a.shape} x {b.shape}")
    f32 = epilogue in (EPI_RESID_F32, EPI_STORE_F32)
    po, ldo = _mat(out, torch.float32 if f32 else torch.bfloat16, "out")
    if tuple(out.shape) != (M, N):
        raise ValueError(f"gemm: out shape {tuple(out.shape)} != {(M, N)}")
    po2 = ldo2 = 0
    if out2 is not None:
        po2, ldo2 = _mat(out2, torch.bfloat16, "out2")
    pax = ldax = 0
    if aux is not None:
        pax, ldax = _mat(aux, torch.float32 if epilogue == EPI_RESID_F32 else torch.bfloat16, "aux")
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != N or not bias.is_contiguous()):
        raise ValueError("gemm: bias must be contiguous fp32 [N]")
    _lib.check(_lib.load().peft_gemm_bf16_rs(pa, lda, pb, ldb, M, N, K, _p(bias), po, ldo, po2, ldo2, pax, ldax, epilogue,
                                             _rowscale(rowscale, M), _stream()), "peft_gemm_bf16")
    return out


def _gemm_f32(a, b, out, bias, epilogue, out2, aux, rowscale=None):
    f32 = torch.float32
    pa, lda = _mat(a, f32, "a")
    pb, ldb = _mat(b, f32, "b")
    M, K = a.shape
    N = b.shape[0]
    if b.shape[1] != K or tuple(out.shape) != (M, N):
        raise ValueError(f"gemm(fp32): shape mismatch {tuple(a.shape)} x {tuple(b.shape)} -> {tuple(out.shape)}")
    po, ldo = _mat(out, f32, "out")
    po2, ldo2 = _mat(out2, f32, "out2") if out2 is not None else (0, 0)
    pax, ldax = _mat(aux, f32, "aux") if aux is not None else (0, 0)
    if bias is not None and (bias.dtype != f32 or bias.numel() != N or not bias.is_contiguous()):
        raise ValueError("gemm: bias must be contiguous fp32 [N]")
    _lib.check(_lib.load().peft_gemm_f32_rs(pa, lda, pb, ldb, M, N, K, _p(bias), po, ldo, po2, ldo2, pax, ldax, epilogue,
                                            _rowscale(rowscale, M), _stream()), "peft_gemm_f32")
    return out


def add(a, b, out):
    """out = a + b, contiguous fp32 (fp32 mode only: the bf16 path fuses this add into layernorm_fwd_add)."""
    for t in (a, b, out):
        if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != out.numel():
            raise ValueError("add: contiguous CUDA fp32 tensors of one size required")
    _lib.check(_lib.load().peft_add_f32(a.data_ptr(), b.data_ptr(), out.data_ptr(), out.numel(), _stream()), "peft_add_f32")
    return out


def layernorm_fwd(x, gamma, beta, y, mean=None, rstd=None, eps=1e-6):
    px, ldx = _mat(x, None, "x")
    py, ldy = _mat(y, None, "y")
    M, D = x.shape
    _lib.check(_lib.load().peft_layernorm_fwd(px, ldx, int(x.dtype == torch.bfloat16), _p(gamma), _p(beta), py, ldy,
                                              int(y.dtype == torch.float32), _p(mean), _p(rstd), M, D, float(eps),
                                              _stream()), "peft_layernorm_fwd")
    return y


def layernorm_fwd_add(x, gamma, beta, y, mean, rstd, add_in, add_out, eps=1e-5):
    """y(bf16) = LN(x fp32); add_out(fp32) = x + add_in(fp32).  fp32 y (verification mode): two unfused launches."""
    if y.dtype == torch.float32:
        layernorm_fwd(x, gamma, beta, y, mean, rstd, eps)
        add(x, add_in, add_out)
        return y
    px, ldx = _mat(x, torch.float32, "x")
    py, ldy = _mat(y, torch.bfloat16, "y")
    pai, ldai = _mat(add_in, torch.float32, "add_in")
    pao, ldao = _mat(add_out, torch.float32, "add_out")
    M, D = x.shape
    _lib.check(_lib.load().peft_layernorm_fwd_add(px, ldx, _p(gamma), _p(beta), py, ldy, _p(mean), _p(rstd), pai, ldai, pao,
                                                  ldao, M, D, float(eps), _stream()), "peft_layernorm_fwd_add")
    return y


def layernorm_bwd(dy, x, mean, rstd, gamma, g_in, g_out, dgamma=None, dbeta=None):
    pdy, lddy = _mat(dy, None, "dy")
    px, ldx = _mat(x, None, "x")
    pgo, ldgo = _mat(g_out, torch.bfloat16, "g_out")
    pgi = ldgi = 0
    if g_in is not None:
        pgi, ldgi = _mat(g_in, torch.bfloat16, "g_in")
    M, D = x.shape
    _lib.check(_lib.load().peft_layernorm_bwd(pdy, lddy, int(dy.dtype == torch.float32), px, ldx,
                                              int(x.dtype == torch.bfloat16), _p(mean), _p(rstd), _p(gamma), pgi, ldgi,
                                              pgo, ldgo, _p(dgamma), _p(dbeta), M, D, _stream()), "peft_layernorm_bwd")
    return g_out


def layernorm_bwd_f32g(dy, x, mean, rstd, gamma, g_in, g_out, g_copy=None, dgamma=None, dbeta=None):
    """fp32 residual-gradient stream: g_out(fp32) = g_in(fp32, may alias) + LN'(dy); g_copy = bf16(g_out)."""
    pdy, lddy = _mat(dy, None, "dy")
    px, ldx = _mat(x, None, "x")
    pgo, ldgo = _mat(g_out, torch.float32, "g_out")
    pgi = ldgi = 0
    if g_in is not None:
        pgi, ldgi = _mat(g_in, torch.float32, "g_in")
    pgc = ldgc = 0
    if g_copy is not None and g_copy.dtype == torch.float32:
        if g_copy.data_ptr() != g_out.data_ptr():      # fp32 mode: the "copy" IS the fp32 stream
            raise ValueError("layernorm_bwd_f32g: an fp32 g_copy must alias g_out")
        g_copy = None
    if g_copy is not None:
        pgc, ldgc = _mat(g_copy, torch.bfloat16, "g_copy")
    M, D = x.shape
    _lib.check(_lib.load().peft_layernorm_bwd_f32g(pdy, lddy, int(dy.dtype == torch.float32), px, ldx,
                                                   int(x.dtype == torch.bfloat16), _p(mean), _p(rstd), _p(gamma), pgi,
                                                   ldgi, pgo, ldgo, pgc, ldgc, _p(dgamma), _p(dbeta), M, D, _stream()),
               "peft_layernorm_bwd_f32g")
    return g_out


def attention_fwd(qkv, o, lse2, B, H, L, scale):
    if qkv.dtype == torch.float32:
        pq, ldq = _mat(qkv, torch.float32, "qkv")
        po, ldo = _mat(o, torch.float32, "o")
        _lib.check(_lib.load().peft_attention_fwd_f32(pq, ldq, po, ldo, _p(lse2), B, H, L, float(scale), _stream()),
                   "peft_attention_fwd_f32")
        return o
    pq, ldq = _mat(qkv, torch.bfloat16, "qkv")
    po, ldo = _mat(o, torch.bfloat16, "o")
    _lib.check(_lib.load().peft_attention_fwd(pq, ldq, po, ldo, _p(lse2), B, H, L, float(scale), _stream()),
               "peft_attention_fwd")
    return o


def attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, scale):
    if qkv.dtype == torch.float32:
        f32 = torch.float32
        (pq, ldq), (po, ldo), (pg, ldg), (pd, ldd) = _mat(qkv, f32, "qkv"), _mat(o, f32, "o"), _mat(d_o, f32, "d_o"), _mat(dqkv, f32, "dqkv")
        _lib.check(_lib.load().peft_attention_bwd_f32(pq, ldq, po, ldo, pg, ldg, _p(lse2), _p(delta), pd, ldd, B, H, L,
                                                      float(scale), _stream()), "peft_attention_bwd_f32")
        return dqkv
    pq, ldq = _mat(qkv, torch.bfloat16, "qkv")
    po, ldo = _mat(o, torch.bfloat16, "o")
    pg, ldg = _mat(d_o, torch.bfloat16, "d_o")
    pd, ldd = _mat(dqkv, torch.bfloat16, "dqkv")
    _lib.check(_lib.load().peft_attention_bwd(pq, ldq, po, ldo, pg, ldg, _p(lse2), _p(delta), pd, ldd, B, H, L,
                                              float(scale), _stream()), "peft_attention_bwd")
    return dqkv


def im2col_patches(images, patches, patch):
    B, C, Hh, Ww = images.shape
    if C != 3 or Hh != Ww or images.dtype != torch.float32 or not images.is_contiguous() or not images.is_cuda:
        raise ValueError("im2col_patches: images must be contiguous CUDA fp32 [B,3,S,S]")
    fn = _lib.load().peft_im2col_patches_f32 if patches.dtype == torch.float32 else _lib.load().peft_im2col_patches
    _lib.check(fn(images.data_ptr(), patches.data_ptr(), patches.stride(0), B, Hh, patch, _stream()), "peft_im2col_patches")
    return patches


def assemble_tokens(patch_embed, cls, pos, x, B, L, D):
    fn = _lib.load().peft_assemble_tokens_f32 if patch_embed.dtype == torch.float32 else _lib.load().peft_assemble_tokens
    _lib.check(fn(patch_embed.data_ptr(), cls.data_ptr(), pos.data_ptr(), x.data_ptr(), B, L, D, _stream()), "peft_assemble_tokens")
    return x


def tn_reduce(x, y, out):
    """out[I,J] += x[M,I]^T @ y[M,J]  (fp32 out, bf16 x / y; fp32 x / y select the verification kernel, any I, J)."""
    if x.dtype == torch.float32:
        (px, ldx), (py, ldy), (po, ldo) = _mat(x, torch.float32, "x"), _mat(y, torch.float32, "y"), _mat(out, torch.float32, "out")
        if tuple(out.shape) != (x.shape[1], y.shape[1]) or y.shape[0] != x.shape[0]:
            raise ValueError("tn_reduce: shape mismatch")
        _lib.check(_lib.load().peft_tn_reduce_f32(px, ldx, x.shape[1], py, ldy, y.shape[1], po, ldo, x.shape[0], _stream()),
                   "peft_tn_reduce_f32")
        return out
    px, ldx = _mat(x, torch.bfloat16, "x")
    py, ldy = _mat(y, torch.bfloat16, "y")
    po, ldo = _mat(out, torch.float32, "out")
    M, I = x.shape
    J = y.shape[1]
    if tuple(out.shape) != (I, J) or y.shape[0] != M:
        raise ValueError("tn_reduce: shape mismatch")
    _lib.check(_lib.load().peft_tn_reduce(px, ldx, I, py, ldy, J, po, ldo, M, _stream()), "peft_tn_reduce")
    return out


def clip_delta(qkv, delta, B, H, L, mult_q, mult_v, backward=False):
    """CLIP-dialect raw-reshape placement of the q/v update, see peft_clip_delta."""
    dt = qkv.dtype if qkv.dtype == torch.float32 else torch.bfloat16
    pq, ldq = _mat(qkv, dt, "qkv")
    pd, ldd = _mat(delta, dt, "delta")
    if qkv.shape[0] != B * L or delta.shape[0] != B * L or delta.shape[1] != 2 * H * 64:
        raise ValueError("clip_delta: shape mismatch")
    fn = _lib.load().peft_clip_delta_f32 if dt == torch.float32 else _lib.load().peft_clip_delta
    _lib.check(fn(pq, ldq, pd, ldd, B, H, L, float(mult_q), float(mult_v), int(backward), _stream()), "peft_clip_delta")
    return delta if backward else qkv


def gemm_tn(x, y, out):
    """out[M, N] (fp32, contiguous) = x[T, M]^T @ y[T, N] (bf16 token-major operands), see peft_gemm_tn_bf16."""
    px, ldx = _mat(x, torch.bfloat16, "x")
    py, ldy = _mat(y, torch.bfloat16, "y")
    po, ldo = _mat(out, torch.float32, "out")
    T, M = x.shape
    N = y.shape[1]
    if y.shape[0] != T or tuple(out.shape) != (M, N):
        raise ValueError(f"gemm_tn: shape mismatch {tuple(x.shape)}^T x {tuple(y.shape)} -> {tuple(out.shape)}")
    _lib.check(_lib.load().peft_gemm_tn_bf16(px, ldx, py, ldy, M, N, T, po, ldo, _stream()), "peft_gemm_tn_bf16")
    return out


def colsum(x, out):
    """out[c] += sum_r x[r, c] (bf16 x, fp32 out [cols], caller-zeroed)."""
    px, ldx = _mat(x, torch.bfloat16, "x")
    if out.dtype != torch.float32 or out.numel() != x.shape[1] or not out.is_contiguous() or not out.is_cuda:
        raise ValueError("colsum: out must be a contiguous CUDA fp32 vector of length cols")
    _lib.check(_lib.load().peft_colsum_bf16(px, ldx, x.shape[0], x.shape[1], out.data_ptr(), _stream()), "peft_colsum_bf16")
    return out


def transpose(x, out, colsum=None):
    """out[c, r] = x[r, c] (bf16); with `colsum` (fp32 [cols], caller-zeroed) also colsum[c] += sum_r x[r, c]."""
    px, ldx = _mat(x, torch.bfloat16, "x")
    po, ldo = _mat(out, torch.bfloat16, "out")
    R, Cc = x.shape
    if tuple(out.shape) != (Cc, R):
        raise ValueError("transpose: out must be [cols, rows]")
    if colsum is None:
        _lib.check(_lib.load().peft_transpose_bf16(px, ldx, R, Cc, po, ldo, _stream()), "peft_transpose_bf16")
    else:
        if colsum.dtype != torch.float32 or colsum.numel() != Cc or not colsum.is_contiguous() or not colsum.is_cuda:
            raise ValueError("transpose: colsum must be a contiguous CUDA fp32 vector of length cols")
        _lib.check(_lib.load().peft_transpose_colsum_bf16(px, ldx, R, Cc, po, ldo, colsum.data_ptr(), _stream()),
                   "peft_transpose_colsum_bf16")
    return out


def phm_fold(rule, x, out, backward=False):
    """forward: out[2, n*n] = x[2, n] @ rule.view(n, n*n); backward: out[2, n] = x[2, n*n] @ rule.view(n, n*n)^T."""
    n = rule.shape[0]
    for t in (rule, x, out):
        if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
            raise RuntimeError("phm_fold: contiguous CUDA fp32 tensors required (no CPU fallback)")
    _lib.check(_lib.load().peft_phm_fold(rule.data_ptr(), x.data_ptr(), out.data_ptr(), n, int(backward), _stream()),
               "peft_phm_fold")
    return out
