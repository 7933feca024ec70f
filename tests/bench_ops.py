"""Micro-benchmarks (not a test): CUDA-event timing of individual C-ABI ops at the headline shapes.

The *_stats / *_phases / gemm_dbg entries read in-kernel counters that exist only in a diagnostics build of the library:
`make -C <package>/csrc clean && make -C <package>/csrc DIAG=1` (the production build compiles them out).

    python tests/bench_ops.py attn|ln|reduce|all
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from peft_b200 import _lib, ops


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def bench_attn(B=256, H=12, L=197):
    D = H * 64
    g = torch.Generator(device="cuda").manual_seed(0)
    qkv = torch.randn(B * L, 3 * D, device="cuda", generator=g).to(torch.bfloat16)
    o = torch.empty(B * L, D, device="cuda", dtype=torch.bfloat16)
    d_o = torch.randn(B * L, D, device="cuda", generator=g).to(torch.bfloat16)
    dqkv = torch.zeros(B * L, 3 * D, device="cuda", dtype=torch.bfloat16)
    lse2 = torch.empty(B * H * L, device="cuda")
    delta = torch.empty(B * H * L, device="cuda")
    scale = 64 ** -0.5
    flops_f = 4.0 * B * H * L * L * 64
    for algo in (2, 3):          # 1 = mma.sync kernels, 2 = tcgen05 with the unpipelined backward, 3 = pipelined backward (default)
        _lib.check(_lib.load().peft_set_attention_algo(algo))
        tf = timeit(lambda: ops.attention_fwd(qkv, o, lse2, B, H, L, scale))
        tb = timeit(lambda: ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, scale))
        print(f"attention B={B} H={H} L={L} algo={algo}: fwd {tf:8.1f} us ({flops_f / tf / 1e6:6.1f} TFLOP/s)  "
              f"bwd {tb:8.1f} us ({2.5 * flops_f / tb / 1e6:6.1f} TFLOP/s)", flush=True)
    _lib.load().peft_set_attention_algo(0)


def bench_ln(M=50432, D=768):
    x = torch.randn(M, D, device="cuda")
    gamma = torch.ones(D, device="cuda")
    beta = torch.zeros(D, device="cuda")
    y = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    mean = torch.empty(M, device="cuda")
    rstd = torch.empty(M, device="cuda")
    t = timeit(lambda: ops.layernorm_fwd(x, gamma, beta, y, mean, rstd, 1e-6))
    print(f"ln_fwd  M={M} D={D}: {t:7.1f} us  {(M * D * 6) / t / 1e3:7.1f} GB/s")
    dy = torch.randn(M, D, device="cuda").to(torch.bfloat16)
    gi = torch.randn(M, D, device="cuda").to(torch.bfloat16)
    go = torch.empty_like(gi)
    t = timeit(lambda: ops.layernorm_bwd(dy, x, mean, rstd, gamma, gi, go))
    print(f"ln_bwd  M={M} D={D}: {t:7.1f} us  {(M * D * 10) / t / 1e3:7.1f} GB/s")
    g32 = torch.randn(M, D, device="cuda")
    gb = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    t = timeit(lambda: ops.layernorm_bwd_f32g(dy, x, mean, rstd, gamma, g32, g32, gb))
    print(f"ln_bwd_f32g (engine form: bf16 dy, fp32 x, fp32 g in place + bf16 copy) M={M} D={D}: {t:7.1f} us  {(M * D * 16) / t / 1e3:7.1f} GB/s")


def bench_ln_pg(D=768):
    """LayerNorm backward with dgamma / dbeta (full mode, Houlsby adapter norm) at the C5 and C4 row counts."""
    for M in (12800, 32896):
        x = torch.randn(M, D, device="cuda")
        dy = torch.randn(M, D, device="cuda").to(torch.bfloat16)
        g = torch.randn(M, D, device="cuda")
        gb = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
        gamma = torch.randn(D, device="cuda")
        mean, rstd = x.mean(1), 1.0 / x.std(1)
        dg, db = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
        t = timeit(lambda: ops.layernorm_bwd_f32g(dy, x, mean, rstd, gamma, g, g, gb, dg, db), iters=20, warm=3)
        byt = M * D * (2 + 4 + 4 + 4 + 2)
        print(f"ln_bwd +dgamma/dbeta M={M} D={D}: {t:7.1f} us  {byt / t / 1e3:7.1f} GB/s", flush=True)


def bench_reduce(M=50432):
    wide = torch.randn(M, 800, device="cuda").to(torch.bfloat16)
    y = torch.randn(M, 2304, device="cuda").to(torch.bfloat16)
    out = torch.zeros(32, 2304, device="cuda")
    t = timeit(lambda: ops.tn_reduce(wide[:, 768:800], y, out))
    print(f"tn_reduce I=32 J=2304 M={M}: {t:7.1f} us  {(M * (2304 + 32) * 2) / t / 1e3:7.1f} GB/s")
    out2 = torch.zeros(32, 768, device="cuda")
    t = timeit(lambda: ops.tn_reduce(y[:, :32], wide[:, :768], out2))
    print(f"tn_reduce I=32 J=768  M={M}: {t:7.1f} us  {(M * (768 + 32) * 2) / t / 1e3:7.1f} GB/s")


def bench_gemm(M=50432, iters=20, only=None, warm=2):
    cases = [("qkv   [x|p][W|Up]^T+b", 2304, 800, ops.EPI_STORE), ("proj  +resid f32", 768, 768, ops.EPI_RESID_F32),
             ("fc1   +GELU (h, act)", 3072, 768, ops.EPI_GELU), ("fc2   +resid f32", 768, 3072, ops.EPI_RESID_F32),
             ("fc2^T *gelu'(h)", 3072, 768, ops.EPI_DGELU), ("fc1^T store", 768, 3072, ops.EPI_STORE),
             ("proj^T store", 768, 768, ops.EPI_STORE), ("qkv^T [dqkv|dp]", 768, 2336, ops.EPI_STORE)]
    if only is not None:
        cases = [cases[i] for i in only]
    for name, N, K, epi in cases:
        a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        b = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
        bias = torch.randn(N, device="cuda")
        f32 = epi == ops.EPI_RESID_F32
        out = torch.empty(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
        out2 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16) if epi == ops.EPI_GELU else None
        aux = None
        if f32:
            aux = torch.randn(M, N, device="cuda")
        elif epi == ops.EPI_DGELU:
            aux = torch.randn(M, N, device="cuda").to(torch.bfloat16)
        t = timeit(lambda: ops.gemm(a, b, out, bias=bias, epilogue=epi, out2=out2, aux=aux), iters=iters, warm=warm)
        print(f"gemm {name:24s} M={M} N={N} K={K}: {t:7.1f} us  {2.0 * M * N * K / t / 1e6:7.1f} TFLOP/s", flush=True)


def gemm_stats(M=50432):
    """PEFT_GEMM_STATS=1: where the three roles of the 2-CTA GEMM wait (cycles, mean over CTAs) for each C3 product."""
    import ctypes
    cases = [("qkv store", 2304, 784, ops.EPI_STORE), ("proj +resid f32", 768, 768, ops.EPI_RESID_F32),
             ("fc1 +GELU", 3072, 768, ops.EPI_GELU), ("fc2 +resid f32", 768, 3072, ops.EPI_RESID_F32),
             ("fc2^T *gelu'", 3072, 768, ops.EPI_DGELU), ("fc1^T store", 768, 3072, ops.EPI_STORE)]
    buf = (ctypes.c_ulonglong * (148 * 8))()
    for name, N, K, epi in cases:
        a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        b = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
        bias = torch.randn(N, device="cuda")
        f32 = epi == ops.EPI_RESID_F32
        out = torch.empty(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
        out2 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16) if epi == ops.EPI_GELU else None
        aux = torch.randn(M, N, device="cuda") if f32 else (torch.randn(M, N, device="cuda").to(torch.bfloat16) if epi == ops.EPI_DGELU else None)
        run = lambda: ops.gemm(a, b, out, bias=bias, epilogue=epi, out2=out2, aux=aux)
        t = timeit(run, iters=5, warm=2)
        run()
        _lib.check(_lib.load().peft_debug_gemm_stats(buf, 148))
        v = torch.tensor(list(buf), dtype=torch.float64).view(148, 8)
        lead = v[0::2]
        tot = float(lead[:, 4].mean())
        print(f"{name:18s} M={M} N={N} K={K}: {t:7.1f} us | issuer total {tot:9.0f} clk: wait-full {100 * float(lead[:, 2].mean()) / tot:5.1f}% "
              f"wait-tmem-empty {100 * float(lead[:, 3].mean()) / tot:5.1f}% | producer wait-empty {100 * float((v[:, 0] / v[:, 1]).mean()):5.1f}% | "
              f"epilogue wait-acc {100 * float((v[:, 5] / v[:, 6]).mean()):5.1f}% | clk/kblock {tot / (((M + 255) // 256) * ((N + 255) // 256) / 74.0 * ((K + 63) // 64)):6.0f}", flush=True)


def gemm_scaling_probe():
    """Per-UMMA pace of the 2-CTA kernel as a function of how many SM pairs are busy: one 256 x 256 tile per cluster with a
    long K (pure MMA when PEFT_GEMM_DBG=12).  Separates a per-SM limit from a chip-level (power / fabric) limit."""
    import ctypes
    _lib.check(_lib.load().peft_set_gemm_algo(2))
    buf = (ctypes.c_ulonglong * (148 * 8))()
    K = 16384
    for clusters in (1, 2, 8, 24, 48, 74):
        M, N = 256 * clusters, 256
        a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        b = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        run = lambda: ops.gemm(a, b, out)
        t = timeit(run, iters=5, warm=2)
        run()
        _lib.check(_lib.load().peft_debug_gemm_stats(buf, 148))
        v = torch.tensor(list(buf), dtype=torch.float64).view(148, 8)[0:2 * clusters:2]
        tot = float(v[:, 4].mean())
        print(f"clusters={clusters:3d} K={K}: {t:7.1f} us  issuer {tot:9.0f} clk  clk/kblock {tot / (K // 64):6.0f}  wait-full {100 * float(v[:, 2].mean()) / tot:5.1f}%  "
              f"=> {2.0 * M * N * K / t / 1e6:7.1f} TFLOP/s, {2.0 * M * N * K / t / 1e6 / clusters:6.2f} per pair", flush=True)
    _lib.check(_lib.load().peft_set_gemm_algo(0))


def graph_time(fn, reps=20, iters=10):
    """us per call of `fn` when `reps` calls are replayed from one CUDA graph (no Python / launch-API time)."""
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        fn(); fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(reps):
                fn()
        g.replay()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (iters * reps) * 1e3


def bench_fixed_costs():
    """Per-launch floor of each kernel family inside a CUDA graph (tiny problems) and the B = 32 layer kernels (M = 6304)."""
    def gemm_case(M, N, K, epi=ops.EPI_STORE, algo=0):
        a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        b = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
        bias = torch.randn(N, device="cuda")
        f32 = epi == ops.EPI_RESID_F32
        out = torch.empty(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
        out2 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16) if epi == ops.EPI_GELU else None
        aux = torch.randn(M, N, device="cuda") if f32 else (torch.randn(M, N, device="cuda").to(torch.bfloat16) if epi == ops.EPI_DGELU else None)
        _lib.check(_lib.load().peft_set_gemm_algo(algo))
        t = graph_time(lambda: ops.gemm(a, b, out, bias=bias, epilogue=epi, out2=out2, aux=aux))
        _lib.check(_lib.load().peft_set_gemm_algo(0))
        return t
    print(f"gemm2 floor  (256x256x64, forced 2-CTA): {gemm_case(256, 256, 64, algo=2):6.1f} us")
    print(f"gemm2 1 wave (18944x256x64 = 74 tiles) : {gemm_case(18944, 256, 64, algo=2):6.1f} us")
    print(f"gemm1 floor  (128x128x64, 1-CTA)       : {gemm_case(128, 128, 64, algo=1):6.1f} us")
    M = 6304
    for name, N, K, epi in [("qkv", 2304, 784, ops.EPI_STORE), ("proj+res", 768, 768, ops.EPI_RESID_F32), ("fc1+gelu", 3072, 768, ops.EPI_GELU),
                            ("fc2+res", 768, 3072, ops.EPI_RESID_F32), ("fc2T*dgelu", 3072, 768, ops.EPI_DGELU), ("fc1T", 768, 3072, ops.EPI_STORE),
                            ("projT", 768, 768, ops.EPI_STORE), ("qkvT", 768, 2320, ops.EPI_STORE), ("p = x Dn^T", 32, 768, ops.EPI_STORE)]:
        t = gemm_case(M, N, K, epi)
        print(f"gemm {name:12s} M={M} N={N} K={K}: {t:6.1f} us  {2.0 * M * N * K / t / 1e6:7.1f} TFLOP/s", flush=True)
    x = torch.randn(M, 768, device="cuda"); gamma = torch.ones(768, device="cuda"); beta = torch.zeros(768, device="cuda")
    y = torch.empty(M, 768, device="cuda", dtype=torch.bfloat16); mean = torch.empty(M, device="cuda"); rstd = torch.empty(M, device="cuda")
    print(f"ln_fwd M={M}: {graph_time(lambda: ops.layernorm_fwd(x, gamma, beta, y, mean, rstd, 1e-6)):6.1f} us")
    dy = torch.randn(M, 768, device="cuda").to(torch.bfloat16); g = torch.randn(M, 768, device="cuda"); gb = torch.empty(M, 768, device="cuda", dtype=torch.bfloat16)
    print(f"ln_bwd_f32g M={M}: {graph_time(lambda: ops.layernorm_bwd_f32g(dy, x, mean, rstd, gamma, g, g, gb)):6.1f} us")
    wide = torch.randn(M, 800, device="cuda").to(torch.bfloat16); yy = torch.randn(M, 2304, device="cuda").to(torch.bfloat16)
    o1 = torch.zeros(32, 2304, device="cuda"); o2 = torch.zeros(32, 768, device="cuda")
    print(f"tn_reduce 32x2304 M={M}: {graph_time(lambda: ops.tn_reduce(wide[:, 768:800], yy, o1)):6.1f} us")
    print(f"tn_reduce 32x768  M={M}: {graph_time(lambda: ops.tn_reduce(yy[:, :32], wide[:, :768], o2)):6.1f} us")
    B, H, L = 32, 12, 197
    qkv = torch.randn(B * L, 2304, device="cuda").to(torch.bfloat16); o = torch.empty(B * L, 768, device="cuda", dtype=torch.bfloat16)
    d_o = torch.randn(B * L, 768, device="cuda").to(torch.bfloat16); dqkv = torch.zeros(B * L, 2304, device="cuda", dtype=torch.bfloat16)
    lse2 = torch.empty(B * H * L, device="cuda"); delta = torch.empty(B * H * L, device="cuda")
    print(f"attn fwd B={B}: {graph_time(lambda: ops.attention_fwd(qkv, o, lse2, B, H, L, 0.125)):6.1f} us")
    print(f"attn bwd B={B}: {graph_time(lambda: ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, 0.125)):6.1f} us")


def bench_gemm_dw(tokens=12800, iters=20):
    """Weight-gradient products of the full mode (C5): out[N_out, K_in] = dY^T X, contraction over the tokens."""
    for name, No, Ki in [("dW qkv", 2304, 768), ("dW proj", 768, 768), ("dW fc1", 3072, 768), ("dW fc2", 768, 3072)]:
        a = torch.randn(No, tokens, device="cuda").to(torch.bfloat16)
        b = torch.randn(Ki, tokens, device="cuda").to(torch.bfloat16)
        out = torch.empty(No, Ki, device="cuda")
        t = timeit(lambda: ops.gemm(a, b, out, epilogue=ops.EPI_STORE_F32), iters=iters, warm=2)
        print(f"gemm {name:10s} M={No} N={Ki} K={tokens}: {t:7.1f} us  {2.0 * No * Ki * tokens / t / 1e6:7.1f} TFLOP/s", flush=True)
        dy = torch.randn(tokens, No, device="cuda").to(torch.bfloat16)
        xx = torch.randn(tokens, Ki, device="cuda").to(torch.bfloat16)
        t = timeit(lambda: ops.gemm_tn(dy, xx, out), iters=iters, warm=2)
        print(f"gemm_tn {name:7s} M={No} N={Ki} T={tokens}: {t:7.1f} us  {2.0 * No * Ki * tokens / t / 1e6:7.1f} TFLOP/s  (no transposes)", flush=True)
        cs = torch.zeros(No, device="cuda")
        t = timeit(lambda: ops.colsum(dy, cs), iters=iters, warm=2)
        print(f"     colsum [{tokens} x {No}] bf16: {t:7.1f} us  {2.0 * tokens * No / t / 1e3:7.1f} GB/s", flush=True)
        x = torch.randn(tokens, No, device="cuda").to(torch.bfloat16)
        xt = torch.empty(No, tokens, device="cuda", dtype=torch.bfloat16)
        t = timeit(lambda: ops.transpose(x, xt), iters=iters, warm=2)
        print(f"     transpose [{tokens} x {No}] bf16: {t:7.1f} us  {4.0 * tokens * No / t / 1e3:7.1f} GB/s", flush=True)


def bench_phm_fold(n=768, peak_gbs=6552.0):
    """Large-n KAdaptation fold: one streaming pass over phm_rule (n^3 fp32 = 1.81 GB at n = 768) per direction."""
    rule = torch.randn(n, n, n, device="cuda")
    s = torch.randn(2, n, device="cuda")
    H = torch.empty(2, n, n, device="cuda")
    dH = torch.randn(2, n, n, device="cuda")
    ds = torch.empty(2, n, device="cuda")
    byt = 4.0 * n ** 3
    tf = timeit(lambda: ops.phm_fold(rule, s, H), iters=10, warm=2)
    tb = timeit(lambda: ops.phm_fold(rule, dH, ds, backward=True), iters=10, warm=2)
    print(f"phm_fold n={n}: fwd {tf:7.1f} us {byt / tf / 1e3:7.1f} GB/s ({byt / tf / 1e3 / peak_gbs:.2f} of copy peak)  "
          f"bwd {tb:7.1f} us {byt / tb / 1e3:7.1f} GB/s ({byt / tb / 1e3 / peak_gbs:.2f})", flush=True)


def vit_b32_shapes(D=768):
    shapes = [(D, 3, 32, 32), (D,), (50, D), (D,), (D,)]
    for _ in range(12):
        shapes += [(D,), (D,), (3 * D, D), (3 * D,), (D, D), (D,), (D,), (D,), (4 * D, D), (4 * D,), (D, 4 * D), (D,)]
    return shapes + [(D,), (D,), (D, 512)]


def bench_fastfood(d=1000, iters=10, peak_gbs=6552.0):
    """C5: Fastfood projection of the whole ViT-B/32 visual parameter set (152 tensors) and its backward."""
    import numpy as np
    from peft_b200.fastfood import FastfoodPlan
    shapes = vit_b32_shapes()
    gn = np.random.RandomState(0)
    params = []
    for s in shapes:
        DD = int(np.prod(s))
        LL = 2 ** int(np.ceil(np.log(DD) / np.log(2)))
        BB = torch.randint(0, 2, (LL,)).float() * 2 - 1
        Pi = torch.from_numpy(gn.permutation(LL)).long()
        GG = torch.randn(LL)
        params.append([BB, Pi, GG, torch.sqrt(LL * torch.sum(GG ** 2)), LL])
    plan = FastfoodPlan(shapes, params, d, "cuda")
    V = torch.randn(d, device="cuda") * 0.01
    plan.dtheta.normal_()
    bf, bb = plan.algorithmic_bytes()
    tf = timeit(lambda: plan.project(V), iters=iters, warm=2)
    tb = timeit(lambda: plan.project_backward(plan.dtheta), iters=iters, warm=2)
    print(f"fastfood ViT-B/32 T={plan.T} sumLL={plan.sum_ll / 1e6:.1f}M sumDD={sum(plan.numels) / 1e6:.2f}M d={d}: "
          f"fwd {tf:8.1f} us  {bf / tf / 1e3:7.1f} GB/s algorithmic ({bf / 1e9:.2f} GB, {bf / tf / 1e3 / peak_gbs:.2f} of copy peak)  "
          f"bwd {tb:8.1f} us  {bb / tb / 1e3:7.1f} GB/s ({bb / 1e9:.2f} GB, {bb / tb / 1e3 / peak_gbs:.2f})", flush=True)
    return plan, V


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("attn", "all"):
        bench_attn()
        bench_attn(128, 16, 257)
        bench_attn(256, 12, 50)
    if what in ("gemm", "all"):
        bench_gemm()
    if what == "small":       # per-GPU batch 32 of the fixed-global-batch split at 8 GPUs (M = 6304)
        bench_gemm(6304, iters=50)
        bench_attn(32, 12, 197)
        bench_ln(6304)
        bench_reduce(6304)
    if what == "attn_stats":
        import ctypes
        for (B, H, L) in ((256, 12, 197), (128, 16, 257)):
            D = H * 64
            qkv = torch.randn(B * L, 3 * D, device="cuda").to(torch.bfloat16)
            o = torch.empty(B * L, D, device="cuda", dtype=torch.bfloat16)
            lse2 = torch.empty(B * H * L, device="cuda")
            t = timeit(lambda: ops.attention_fwd(qkv, o, lse2, B, H, L, 0.125), iters=5, warm=2)
            buf = (ctypes.c_ulonglong * (148 * 16))()
            _lib.check(_lib.load().peft_debug_attn_stats(buf, 148))
            v = torch.tensor(list(buf), dtype=torch.float64).view(148, 2, 8)
            tot = v[:, :, 5].mean()
            names = ["wait S", "max pass", "exp pass", "wait O", "epilogue"]
            print(f"attn fwd B={B} H={H} L={L}: {t:.1f} us; softmax warp total {tot:.0f} clk per CTA ({B * H / 148:.1f} heads): " +
                  ", ".join(f"{n} {100 * float(v[:, :, i].mean() / tot):.1f}%" for i, n in enumerate(names)), flush=True)
    if what == "attn_bwd_stats":   # PEFT_ATTN_STATS=1: cycle counters of the pipelined backward (attention3_bwd_sm100.cuh)
        import ctypes
        for (B, H, L) in ((256, 12, 197), (128, 16, 257), (256, 12, 50)):
            D = H * 64
            qkv = torch.randn(B * L, 3 * D, device="cuda").to(torch.bfloat16)
            o = torch.empty(B * L, D, device="cuda", dtype=torch.bfloat16)
            d_o = torch.randn(B * L, D, device="cuda").to(torch.bfloat16)
            dqkv = torch.zeros(B * L, 3 * D, device="cuda", dtype=torch.bfloat16)
            lse2 = torch.empty(B * H * L, device="cuda")
            delta = torch.empty(B * H * L, device="cuda")
            ops.attention_fwd(qkv, o, lse2, B, H, L, 0.125)
            t = timeit(lambda: ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, 0.125), iters=5, warm=2)
            buf = (ctypes.c_ulonglong * (512 * 16))()
            _lib.check(_lib.load().peft_debug_attn_stats(buf, 512))
            v = torch.tensor(list(buf), dtype=torch.float64).view(2, 256, 16)[:, :148]
            tiles = B * H * ((L + 127) // 128) / 148
            print(f"attn bwd B={B} H={H} L={L}: {t:.1f} us both kernels, {tiles:.1f} tiles per CTA", flush=True)
            for k, nm in enumerate(("dQ ", "dKV")):
                m = v[k].mean(0)
                print(f"  {nm} MMA warp: total {m[0]:.0f} clk = {m[0] / tiles:.0f} per tile; in issue_ss (operand waits) {100 * m[1] / m[0]:.1f}%, "
                      f"wait dS {100 * m[2] / m[0]:.1f}%, wait accumulator free {100 * m[3] / m[0]:.1f}%")
                for w in range(2):
                    c = m[4 + 6 * w:10 + 6 * w]
                    print(f"  {nm} warpgroup {w}: total {c[0]:.0f} clk; delta/lse {100 * c[1] / c[0]:.1f}%, wait S/dP {100 * c[2] / c[0]:.1f}%, "
                          f"math {100 * c[3] / c[0]:.1f}%, epilogue {100 * c[4] / c[0]:.1f}%")
    if what == "mma_probe":        # dependent-accumulate latency of tcgen05.mma: clk per instruction vs number of accumulators
        import ctypes
        buf = (ctypes.c_ulonglong * 2)()
        for ts in (0, 1):
            for n in (16, 64, 128, 208, 256):
                row = []
                for nacc in (1, 2, 4):
                    if nacc * n > 256:
                        continue
                    _lib.check(_lib.load().peft_debug_mma_probe(n, nacc, ts, 256, buf))
                    row.append(f"nacc={nacc}: {buf[1] / 256:6.1f} clk/mma (issue {buf[0] / 256:5.1f})")
                print(f"{'TS' if ts else 'SS'} M=128 N={n:3d} K=16 (floor {n / 2:5.1f} clk): " + "; ".join(row), flush=True)
    if what == "gemm_once":        # one launch of each of the 8 products of a C3 layer: the command ncu --set full captures
        bench_gemm(iters=1, warm=0)   # (tools/ncu_traffic.py -> profiles/gemm_traffic.json)
    if what == "other_once":       # one launch of the non-GEMM kernels of a C3 layer (ncu --set full target)
        B, H, L, D = 256, 12, 197, 768
        qkv = torch.randn(B * L, 3 * D, device="cuda").to(torch.bfloat16)
        o = torch.empty(B * L, D, device="cuda", dtype=torch.bfloat16)
        d_o = torch.randn(B * L, D, device="cuda").to(torch.bfloat16)
        dqkv = torch.zeros(B * L, 3 * D, device="cuda", dtype=torch.bfloat16)
        lse2 = torch.empty(B * H * L, device="cuda")
        delta = torch.empty(B * H * L, device="cuda")
        ops.attention_fwd(qkv, o, lse2, B, H, L, 0.125)
        ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, 0.125)
        _lib.check(_lib.load().peft_set_attention_algo(3))
        ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, 0.125)
        _lib.load().peft_set_attention_algo(0)
        wide = torch.randn(B * L, 800, device="cuda").to(torch.bfloat16)
        out = torch.zeros(32, 2304, device="cuda")
        ops.tn_reduce(wide[:, 768:800], dqkv, out)
        out2 = torch.zeros(32, 768, device="cuda")
        ops.tn_reduce(dqkv[:, :32], wide[:, :768], out2)
        torch.cuda.synchronize()
    if what == "fixed":
        bench_fixed_costs()
    if what == "gemm_probe":
        print("PEFT_GEMM_DBG =", os.environ.get("PEFT_GEMM_DBG", "0"), flush=True)
        gemm_scaling_probe()
    if what == "gemm_phases":   # PEFT_GEMM_STATS=1 PEFT_GEMM_DBG=128(+...): cycles of epilogue warp 4 per phase
        import ctypes
        M = 50432
        buf = (ctypes.c_ulonglong * (148 * 8))()
        for name, N, K, epi in [("qkv store", 2304, 784, ops.EPI_STORE), ("fc1 +GELU", 3072, 768, ops.EPI_GELU), ("fc2 +resid f32", 768, 3072, ops.EPI_RESID_F32),
                                ("fc2^T *gelu'", 3072, 768, ops.EPI_DGELU)]:
            a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            b = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
            bias = torch.randn(N, device="cuda")
            f32 = epi == ops.EPI_RESID_F32
            out = torch.empty(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
            out2 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16) if epi == ops.EPI_GELU else None
            aux = torch.randn(M, N, device="cuda") if f32 else (torch.randn(M, N, device="cuda").to(torch.bfloat16) if epi == ops.EPI_DGELU else None)
            if os.environ.get("NOBIAS"):
                bias = None
            run = lambda: ops.gemm(a, b, out, bias=bias, epilogue=epi, out2=out2, aux=aux)
            t = timeit(run, iters=5, warm=2)
            run()
            _lib.check(_lib.load().peft_debug_gemm_stats(buf, 148))
            v = torch.tensor(list(buf), dtype=torch.float64).view(148, 8).mean(0)
            units = ((M + 255) // 256) * ((N + 255) // 256) / 74.0 * (4 if f32 else 2)
            names = ["tmem ld", "bias", "slot/aux", "pack+st+store", "act math"]
            print(f"{name:16s} {t:7.1f} us | epilogue warp total {v[6]:9.0f} clk, wait-acc {100 * v[5] / v[6]:4.1f}% | per unit: " +
                  ", ".join(f"{n} {v[i] / units:6.0f}" for i, n in enumerate(names)), flush=True)
    if what == "gemm_stats":
        gemm_stats(50432)
        gemm_stats(6304)
    if what == "gemm_dbg":    # PEFT_GEMM_DBG diagnostics: plain-store and GELU products only (results are wrong by design)
        print("PEFT_GEMM_DBG =", os.environ.get("PEFT_GEMM_DBG", "0"), flush=True)
        bench_gemm(50432, iters=10, only=(0, 2, 5))
        if os.environ.get("PEFT_GEMM_STATS"):
            gemm_stats(50432)
    if what == "prof":      # one pass of each hot kernel, for ncu captures
        bench_gemm(iters=1)
        B, H, L = 256, 12, 197
        D = H * 64
        qkv = torch.randn(B * L, 3 * D, device="cuda").to(torch.bfloat16)
        o = torch.empty(B * L, D, device="cuda", dtype=torch.bfloat16)
        d_o = torch.randn(B * L, D, device="cuda").to(torch.bfloat16)
        dqkv = torch.zeros(B * L, 3 * D, device="cuda", dtype=torch.bfloat16)
        lse2 = torch.empty(B * H * L, device="cuda")
        delta = torch.empty(B * H * L, device="cuda")
        for _ in range(2):
            ops.attention_fwd(qkv, o, lse2, B, H, L, 0.125)
            ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, 0.125)
        bench_ln()
        bench_reduce()
        torch.cuda.synchronize()
    if what == "prof_once":   # exactly ONE launch of each hot kernel at the C3 shapes, for `ncu --set full -k regex:peft`
        bench_gemm(iters=1, warm=0)
        B, H, L = 256, 12, 197
        D = H * 64
        qkv = torch.randn(B * L, 3 * D, device="cuda").to(torch.bfloat16)
        o = torch.empty(B * L, D, device="cuda", dtype=torch.bfloat16)
        d_o = torch.randn(B * L, D, device="cuda").to(torch.bfloat16)
        dqkv = torch.zeros(B * L, 3 * D, device="cuda", dtype=torch.bfloat16)
        lse2 = torch.empty(B * H * L, device="cuda")
        delta = torch.empty(B * H * L, device="cuda")
        ops.attention_fwd(qkv, o, lse2, B, H, L, 0.125)
        ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, 0.125)
        M = B * L
        x = torch.randn(M, D, device="cuda"); gamma = torch.ones(D, device="cuda"); beta = torch.zeros(D, device="cuda")
        y = torch.empty(M, D, device="cuda", dtype=torch.bfloat16); mean = torch.empty(M, device="cuda"); rstd = torch.empty(M, device="cuda")
        ops.layernorm_fwd(x, gamma, beta, y, mean, rstd, 1e-6)
        g32 = torch.randn(M, D, device="cuda"); gb = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
        ops.layernorm_bwd_f32g(d_o, x, mean, rstd, gamma, g32, g32, gb)
        wide = torch.randn(M, 800, device="cuda").to(torch.bfloat16)
        out = torch.zeros(32, 2304, device="cuda")
        ops.tn_reduce(wide[:, 768:800], dqkv, out)
        out2 = torch.zeros(32, 768, device="cuda")
        ops.tn_reduce(dqkv[:, :32], wide[:, :768], out2)
        torch.cuda.synchronize()
    if what in ("fold", "all"):
        bench_phm_fold()
    if what in ("fastfood", "all"):
        bench_fastfood()
    if what == "ln_pg":
        bench_ln_pg()
    if what == "prof_c5":     # one pass of the full-mode (C5) products and transposes, for ncu
        bench_gemm_dw(12800, iters=1)
    if what == "gemm_c5":
        bench_gemm(12800)
        bench_gemm_dw(12800)
    if what == "prof_fastfood":
        bench_fastfood(iters=1)
        torch.cuda.synchronize()
    if what in ("ln", "all"):
        bench_ln()
    if what in ("reduce", "all"):
        bench_reduce()
