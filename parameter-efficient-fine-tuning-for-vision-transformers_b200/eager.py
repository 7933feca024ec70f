"""Module-level (eager) forwards of the timm-dialect sub-modules — `Mlp`, `Attention`, `Adapter`, `LayerNorm`, `Block` — each ONE
autograd node over the C-ABI kernels, with gradients for the input AND for every weight.

Why they exist: the reference measures intrinsic dimension per block component by wrapping stand-alone sub-modules,
`FastFoodWrap(Attention(...))`, `FastFoodWrap(Mlp(...))`, `FastFoodWrap(Adapter(...))`
(full_shot/main/lib/models/cls_intrinsic_dimension.py:466-481, dispatched at :483-500), and `FastfoodWrap` re-sets the wrapped
parameters as plain tensors before every forward (full_shot/main/intrinsic/fastfood.py:61-77).  The whole-tower engine
(engine.py) is the throughput path; these nodes give the same kernels at sub-module granularity: weights are read through
attribute access on every call, dW comes from the token-major tensor-core product (`peft_gemm_tn_bf16`), biases from the
column-sum kernel.  `model.precision = "fp32"`-style verification: pass fp32 = True (plain fp32 kernels).

Reference arithmetic: Mlp FS/kadaptation.py:57-78; Attention :198-238 (no q/v adapter); Adapter FS/adapter.py:308-400
(LayerNorm -> Linear -> ReLU -> Linear, + residual_input); Block :273-277.  No torch math beyond reshapes / dtype casts / the
residual `+` of Block.
"""
from __future__ import annotations

import torch

from . import ops


def _cdt(fp32):
    return torch.float32 if fp32 else torch.bfloat16


def _dw(dY, X):
    """[N, K] fp32 = dY[T, N]^T X[T, K]"""
    out = torch.empty(dY.shape[1], X.shape[1], device=dY.device, dtype=torch.float32)
    if dY.dtype == torch.float32:
        out.zero_()
        ops.tn_reduce(dY, X, out)
    else:
        ops.gemm_tn(dY, X, out)
    return out


def _colsum(dY):
    n = dY.shape[1]
    if dY.dtype == torch.float32:
        ones = torch.zeros(dY.shape[0], 16, device=dY.device, dtype=torch.float32)
        ones[:, 0] = 1.0
        c16 = torch.zeros(16, n, device=dY.device, dtype=torch.float32)
        ops.tn_reduce(ones, dY, c16)
        return c16[0].contiguous()
    return ops.colsum(dY, torch.zeros(n, device=dY.device, dtype=torch.float32))


def _require_cuda(x, what):
    if not x.is_cuda:
        raise RuntimeError(f"peft_b200 {what} runs on CUDA only (no CPU fallback)")


def _f(t):
    return None if t is None else t.detach().float().contiguous()


class _LNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        _require_cuda(x, "LayerNorm")
        D = x.shape[-1]
        x2 = x.detach().reshape(-1, D).float().contiguous()
        y = torch.empty_like(x2)
        mean = torch.empty(x2.shape[0], device=x.device, dtype=torch.float32)
        rstd = torch.empty_like(mean)
        g = _f(gamma)
        ops.layernorm_fwd(x2, g, _f(beta), y, mean, rstd, eps)
        ctx.save_for_backward(x2, mean, rstd, g)
        ctx.shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, mean, rstd, g = ctx.saved_tensors
        D = x2.shape[1]
        dy2 = dy.reshape(-1, D).float().contiguous()
        dx = torch.empty_like(x2)
        wg = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        dg = torch.zeros(D, device=x2.device, dtype=torch.float32) if wg else None
        db = torch.zeros_like(dg) if wg else None
        ops.layernorm_bwd_f32g(dy2, x2, mean, rstd, g, None, dx, None, dg, db)
        return dx.view(ctx.shape), dg, db, None


def layer_norm(x, gamma, beta, eps):
    return _LNFn.apply(x, gamma, beta, float(eps))


class _MlpFn(torch.autograd.Function):
    """y = fc2(act(fc1(x)));  act: ops.EPI_GELU (timm) | ops.EPI_QUICKGELU (CLIP)."""

    @staticmethod
    def forward(ctx, x, W1, b1, W2, b2, act_code, fp32):
        _require_cuda(x, "Mlp")
        cdt = _cdt(fp32)
        C, F = W1.shape[1], W1.shape[0]
        xb = x.detach().reshape(-1, C).to(cdt).contiguous()
        M = xb.shape[0]
        W1c, W2c = W1.detach().to(cdt).contiguous(), W2.detach().to(cdt).contiguous()
        h = torch.empty(M, F, device=x.device, dtype=cdt)
        act = torch.empty_like(h)
        ops.gemm(xb, W1c, h, bias=_f(b1), epilogue=act_code, out2=act)
        y = torch.empty(M, W2.shape[0], device=x.device, dtype=torch.float32)
        ops.gemm(act, W2c, y, bias=_f(b2), epilogue=ops.EPI_STORE_F32)
        ctx.save_for_backward(xb, W1c, W2c, h, act)
        ctx.meta = (x.shape, act_code, b1 is not None, b2 is not None)
        return y.view(*x.shape[:-1], W2.shape[0])

    @staticmethod
    def backward(ctx, dy):
        xb, W1c, W2c, h, act = ctx.saved_tensors
        shape, act_code, hb1, hb2 = ctx.meta
        cdt = xb.dtype
        dact = {ops.EPI_GELU: ops.EPI_DGELU, ops.EPI_QUICKGELU: ops.EPI_DQUICKGELU, ops.EPI_GELU_TANH: ops.EPI_DGELU_TANH}[act_code]
        dyb = dy.reshape(-1, W2c.shape[0]).to(cdt).contiguous()
        need = ctx.needs_input_grad
        dW2 = _dw(dyb, act) if need[3] else None
        db2 = _colsum(dyb) if hb2 and need[4] else None
        dh = torch.empty_like(h)
        ops.gemm(dyb, W2c.t().contiguous(), dh, epilogue=dact, aux=h)
        dW1 = _dw(dh, xb) if need[1] else None
        db1 = _colsum(dh) if hb1 and need[2] else None
        dx = None
        if need[0]:
            dx = torch.empty(xb.shape[0], xb.shape[1], device=xb.device, dtype=torch.float32)
            ops.gemm(dh, W1c.t().contiguous(), dx, epilogue=ops.EPI_STORE_F32)
            dx = dx.view(shape)
        return dx, dW1, db1, dW2, db2, None, None


class _AttnFn(torch.autograd.Function):
    """y = proj(softmax(q k^T * scale) v) with [q | k | v] = qkv(x), heads of 64 (FS/kadaptation.py:219-238, adapter-free)."""

    @staticmethod
    def forward(ctx, x, Wqkv, bqkv, Wproj, bproj, H, scale, fp32):
        _require_cuda(x, "Attention")
        if x.dim() != 3:
            raise ValueError("Attention expects [B, N, C]")
        cdt = _cdt(fp32)
        B, L, C = x.shape
        xb = x.detach().reshape(B * L, C).to(cdt).contiguous()
        Wq, Wp = Wqkv.detach().to(cdt).contiguous(), Wproj.detach().to(cdt).contiguous()
        qkv = torch.empty(B * L, 3 * C, device=x.device, dtype=cdt)
        ops.gemm(xb, Wq, qkv, bias=_f(bqkv))
        o = torch.empty(B * L, C, device=x.device, dtype=cdt)
        lse2 = torch.empty(B * H * L, device=x.device, dtype=torch.float32)
        ops.attention_fwd(qkv, o, lse2, B, H, L, scale)
        y = torch.empty(B * L, C, device=x.device, dtype=torch.float32)
        ops.gemm(o, Wp, y, bias=_f(bproj), epilogue=ops.EPI_STORE_F32)
        ctx.save_for_backward(xb, Wq, Wp, qkv, o, lse2)
        ctx.meta = (B, L, C, H, scale, bqkv is not None, bproj is not None)
        return y.view(B, L, C)

    @staticmethod
    def backward(ctx, dy):
        xb, Wq, Wp, qkv, o, lse2 = ctx.saved_tensors
        B, L, C, H, scale, hbq, hbp = ctx.meta
        cdt = xb.dtype
        dyb = dy.reshape(B * L, C).to(cdt).contiguous()
        need = ctx.needs_input_grad
        dWp = _dw(dyb, o) if need[3] else None
        dbp = _colsum(dyb) if hbp and need[4] else None
        d_o = torch.empty_like(o)
        ops.gemm(dyb, Wp.t().contiguous(), d_o)
        dqkv = torch.zeros_like(qkv)
        delta = torch.empty(B * H * L, device=xb.device, dtype=torch.float32)
        ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, scale)
        dWq = _dw(dqkv, xb) if need[1] else None
        dbq = _colsum(dqkv) if hbq and need[2] else None
        dx = None
        if need[0]:
            dx = torch.empty(B * L, C, device=xb.device, dtype=torch.float32)
            ops.gemm(dqkv, Wq.t().contiguous(), dx, epilogue=ops.EPI_STORE_F32)
            dx = dx.view(B, L, C)
        return dx, dWq, dbq, dWp, dbp, None, None, None


class _AdapterFn(torch.autograd.Function):
    """out = residual + up(relu(down(LN(x))))      (FS/adapter.py:372-400: adapter_layernorm_option "in", scalar 1)."""

    @staticmethod
    def forward(ctx, x, residual, g, b, Wd, bd, Wu, bu, eps, scale, fp32):
        _require_cuda(x, "Adapter")
        cdt = _cdt(fp32)
        C = x.shape[-1]
        x2 = x.detach().reshape(-1, C).float().contiguous()
        r2 = residual.detach().reshape(-1, C).float().contiguous()
        M = x2.shape[0]
        Wdc, Wuc = Wd.detach().to(cdt).contiguous(), (Wu.detach() * scale).to(cdt).contiguous()
        gam = _f(g)
        z0 = torch.empty(M, C, device=x.device, dtype=cdt)
        mean = torch.empty(M, device=x.device, dtype=torch.float32)
        rstd = torch.empty_like(mean)
        ops.layernorm_fwd(x2, gam, _f(b), z0, mean, rstd, eps)
        zd = torch.empty(M, Wd.shape[0], device=x.device, dtype=cdt)
        ops.gemm(z0, Wdc, zd, bias=_f(bd), epilogue=ops.EPI_RELU)
        out = torch.empty(M, C, device=x.device, dtype=torch.float32)
        bu_s = None if bu is None else (_f(bu) * scale)
        ops.gemm(zd, Wuc, out, bias=bu_s, epilogue=ops.EPI_RESID_F32, aux=r2)
        ctx.save_for_backward(x2, mean, rstd, gam, z0, zd, Wdc, Wuc)
        ctx.meta = (x.shape, scale, bd is not None, bu is not None)
        down = zd.view(*x.shape[:-1], zd.shape[1])              # the reference's second return value (post-ReLU activations)
        ctx.mark_non_differentiable(down)
        return out.view(x.shape), down

    @staticmethod
    def backward(ctx, dout, _ddown):
        x2, mean, rstd, gam, z0, zd, Wdc, Wuc = ctx.saved_tensors
        shape, scale, hbd, hbu = ctx.meta
        cdt = z0.dtype
        C = x2.shape[1]
        d2 = dout.reshape(-1, C).float().contiguous()
        dup = d2.to(cdt)
        need = ctx.needs_input_grad
        dWu = _dw(dup, zd) * scale if need[6] else None
        dbu = _colsum(dup) * scale if hbu and need[7] else None
        dzd = torch.empty_like(zd)
        ops.gemm(dup, Wuc.t().contiguous(), dzd, epilogue=ops.EPI_DRELU, aux=zd)
        dWd = _dw(dzd, z0) if need[4] else None
        dbd = _colsum(dzd) if hbd and need[5] else None
        dz0 = torch.empty_like(z0)
        ops.gemm(dzd, Wdc.t().contiguous(), dz0)
        dx = torch.empty_like(x2)
        wg = need[2] or need[3]
        dg = torch.zeros(C, device=x2.device, dtype=torch.float32) if wg else None
        db = torch.zeros_like(dg) if wg else None
        ops.layernorm_bwd_f32g(dz0, x2, mean, rstd, gam, None, dx, None, dg, db)
        return dx.view(shape), d2.view(shape), dg, db, dWd, dbd, dWu, dbu, None, None, None


def mlp(x, W1, b1, W2, b2, act_code=ops.EPI_GELU, fp32=False):
    return _MlpFn.apply(x, W1, b1, W2, b2, act_code, fp32)


def attention(x, Wqkv, bqkv, Wproj, bproj, num_heads, scale, fp32=False):
    if Wqkv.shape[1] != 64 * num_heads:
        raise NotImplementedError("peft_b200 attention kernels need a head dimension of 64")
    return _AttnFn.apply(x, Wqkv, bqkv, Wproj, bproj, int(num_heads), float(scale), fp32)


def adapter(x, residual, ln_weight, ln_bias, Wd, bd, Wu, bu, eps=1e-5, scale=1.0, fp32=False):
    return _AdapterFn.apply(x, residual, ln_weight, ln_bias, Wd, bd, Wu, bu, float(eps), float(scale), fp32)


class _PatchEmbedFn(torch.autograd.Function):
    """Conv2d(3, D, kernel = stride = P) as im2col + product: [B, 3, S, S] -> [B, (S/P)^2, D]  (FS/kadaptation.py:286-308)."""

    @staticmethod
    def forward(ctx, images, W, b, fp32):
        _require_cuda(images, "PatchEmbed")
        cdt = _cdt(fp32)
        D, P = W.shape[0], W.shape[-1]
        B, S = images.shape[0], images.shape[-1]
        K = 3 * P * P
        Kp = (K + 7) // 8 * 8                                  # TMA row pitch: K padded with zero columns
        npatch = (S // P) ** 2
        patches = torch.zeros(B * npatch, Kp, device=images.device, dtype=cdt)
        ops.im2col_patches(images.detach().float().contiguous(), patches, P)
        Wc = torch.zeros(D, Kp, device=images.device, dtype=cdt)
        Wc[:, :K] = W.detach().reshape(D, K)
        y = torch.empty(B * npatch, D, device=images.device, dtype=torch.float32)
        ops.gemm(patches, Wc, y, bias=_f(b), epilogue=ops.EPI_STORE_F32)
        ctx.save_for_backward(patches)
        ctx.meta = (tuple(W.shape), K, b is not None)
        return y.view(B, npatch, D)

    @staticmethod
    def backward(ctx, dy):
        (patches,) = ctx.saved_tensors
        wshape, K, hb = ctx.meta
        need = ctx.needs_input_grad
        dyb = dy.reshape(-1, wshape[0]).to(patches.dtype).contiguous()
        dW = _dw(dyb, patches)[:, :K].reshape(wshape) if need[1] else None
        db = _colsum(dyb) if hb and need[2] else None
        return None, dW, db, None


def patch_embed(images, W, b, fp32=False):
    return _PatchEmbedFn.apply(images, W, b, fp32)
