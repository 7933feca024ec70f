#!/usr/bin/env python
"""Fine-tuning step throughput (images/s, fwd+bwd+all-reduce+SGD) of the B200-native PEFT-ViT engine.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3|C2|...]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[2], the one the metric is quoted on): ViT-B/16 (L=197, D=768, 12 layers) +
KAdaptation (Kronecker q/v update, n=4, rank 1) + BatchNorm/Linear head, synthetic 224x224 images, bf16 compute,
GLOBAL batch 256 split over the N ranks (per-GPU batch 256 / N: the fixed-global-batch partitioning of SURVEY 8e,
"scaling": "strong"; `--scaling weak` keeps per-GPU batch 256), SGD momentum 0.9.  One "step" = zero grads -> forward ->
CrossEntropy -> backward -> adapter-gradient all-reduce (N > 1) -> optimizer step.
Rank 0 prints ONE JSON line.  `--impl reference` times the CPU oracle (the restatement of the reference's PyTorch path,
pinned to the reference by tests/golden) on the host cores for the same model/metric on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

# FLOP per image, fwd + bwd with a frozen backbone (dX only, no dW): SURVEY.md §8d / BASELINE.md §3


def kernel_source_hash():
    """sha256 over the CUDA sources: ties a committed ncu traffic figure to the build it was captured from."""
    import glob
    import hashlib
    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "parameter-efficient-fine-tuning-for-vision-transformers_b200", "csrc")
    for f in sorted(glob.glob(os.path.join(csrc, "*.cu")) + glob.glob(os.path.join(csrc, "*.cuh"))):
        if os.path.basename(f).startswith("test_"):
            continue
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def ncu_gemm_traffic(cfg_name):
    """Mean DRAM bytes (read + write) per GEMM launch from the committed `ncu --set full` summary of THIS build
    (profiles/gemm_traffic.json, written by tools/ncu_traffic.py); None when the sources changed since the capture."""
    p = os.path.join(ROOT, "profiles", "gemm_traffic.json")
    try:
        with open(p) as f:
            d = json.load(f)
    except Exception:
        return None
    if d.get("source_hash") != kernel_source_hash():
        return None
    return d.get("bytes_per_launch", {}).get(cfg_name)

CONFIGS = {
    "C3": dict(model="ViT-B/16", dialect="timm", adapter="kron", patch=16, width=768, depth=12, heads=12, batch=256,
               gflop_img=71.45 + 0.17, desc="ViT-B/16 + KAdaptation (n=4, rank 1) on q/v"),
    # SURVEY Q4 variant: the COMMITTED KAdaptation hyper-parameters (phm_dim = 768, rank 1): dense fold path, phm_rule frozen
    "C3n768": dict(model="ViT-B/16", dialect="timm", adapter="kron", phm_dim=768, patch=16, width=768, depth=12, heads=12,
                   batch=256, gflop_img=71.45 + 3 * 12 * 197 * 2 * 2 * 768 * 768 / 1e9,
                   desc="ViT-B/16 + KAdaptation with the committed phm_dim=768 (dense fold variant)"),
    "C2": dict(model="ViT-B/16", dialect="timm", adapter="lora", patch=16, width=768, depth=12, heads=12, batch=256,
               gflop_img=71.45 + 0.17, desc="ViT-B/16 + LoRA (r=4, alpha=128) on q/v"),
    "B32": dict(model="ViT-B/32", dialect="timm", adapter="kron", patch=32, width=768, depth=12, heads=12, batch=256,
                gflop_img=17.50 + 0.04, desc="ViT-B/32 + KAdaptation (n=4, rank 1) on q/v"),
    # BASELINE.json configs[0]: the reference's own CPU-runnable case (CLIP dialect: Q1/Q2 delta placement)
    "C1": dict(model="CLIP ViT-B/32", dialect="clip", adapter="kron", patch=32, width=768, depth=12, heads=12, batch=32,
               out_dim=512, gflop_img=17.50 + 0.04, desc="CLIP ViT-B/32 visual + KAdaptation (n=4, rank 1) on q/v"),
    # BASELINE.json configs[3]: CLIP ViT-L/14 (seq 257) + Houlsby / Compacter (bottleneck 64), batch 128
    "C4": dict(model="CLIP ViT-L/14", dialect="clip", adapter="houlsby", patch=14, width=1024, depth=24, heads=16, batch=128,
               out_dim=768, gflop_img=330.2 + 4.85, desc="CLIP ViT-L/14 visual + Houlsby adapter (bottleneck 64)"),
    # BASELINE.json configs[4]: intrinsic-dimension fine-tuning, Fastfood d=1000 over the whole visual tower (dW needed)
    "C5": dict(model="ViT-B/32", dialect="timm", adapter="none", patch=32, width=768, depth=12, heads=12, batch=256,
               gflop_img=26.22, fastfood_d=1000, desc="ViT-B/32 + Fastfood intrinsic dimension d=1000 over all 152 tower tensors"),
    "C4c": dict(model="CLIP ViT-L/14", dialect="clip", adapter="compacter", patch=14, width=1024, depth=24, heads=16, batch=128,
                out_dim=768, gflop_img=330.2 + 4.85, desc="CLIP ViT-L/14 visual + Compacter (PHM n=32/4, bottleneck 64)"),
}


def vit_cfg(c, depth=None):
    from peft_b200.synth import VitCfg
    return VitCfg(dialect=c["dialect"], adapter=c["adapter"], img=224, patch=c["patch"], width=c["width"],
                  depth=depth or c["depth"], heads=c["heads"], out_dim=c.get("out_dim", 512), phm_dim=c.get("phm_dim", 4))


def adapter_kwargs(c):
    if c["adapter"] == "kron":
        return dict(phm_dim=c.get("phm_dim", 4), phm_rank=1)
    if c["adapter"] == "lora":
        return dict(lora_attn_dim=4, lora_attn_alpha=128)
    return dict(bottleneck=64)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(tflops=float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1400.0))),
                    tflops_burst=float(d.get("bf16_tflops", 1590.0)), hbm=float(d.get("hbm_gbs", 6650.0)), src="measured")
    return dict(tflops=1400.0, tflops_burst=1590.0, hbm=6650.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def build_oracle_step(cfg_name, batch):
    from oracle import peft_oracle as O
    from oracle import synth
    c = CONFIGS[cfg_name]
    cfg = vit_cfg(c)
    sd = synth.make_state_dict(cfg, seed=0, grad_test=False)
    head = synth.make_head(cfg)
    names = synth.trainable_names(cfg, sd)
    images, labels = synth.make_batch(cfg, batch)
    if c.get("fastfood_d"):
        import numpy as np
        d = c["fastfood_d"]
        gn, gt = np.random.RandomState(0), torch.Generator().manual_seed(0)
        ff = {k: O.fastfood_vars(v.numel(), gn, gt) for k, v in sd.items()}
        V = torch.zeros(d)
        momV = torch.zeros(d)

        def ff_step():
            V.requires_grad_(True)
            V.grad = None
            sdp = {k: v + O.fastfood_torched(V, v.numel(), ff[k]).view(v.shape) for k, v in sd.items()}
            feat = O.visual_forward(images, sdp, cfg)
            loss = torch.nn.functional.cross_entropy(O.classifier_logits(feat, head, True), labels)
            loss.backward()
            with torch.no_grad():
                momV.mul_(0.9).add_(V.grad)
                V.requires_grad_(False)
                V.sub_(1e-3 * momV)
            return float(loss)
        return ff_step
    mom = {k: torch.zeros_like(sd[k]) for k in names}

    def step():
        logits, loss, grads, _ = O.train_step(images, labels, sd, head, cfg, names)
        with torch.no_grad():   # SGD momentum 0.9 (vitb32_CLIP.yaml:40-43)
            for k in names:
                mom[k].mul_(0.9).add_(grads[k])
                sd[k].sub_(1e-3 * mom[k])
        return float(loss)
    return step


def cpu_reference_rate(cfg_name, budget_s, steps=None, warmup=1):
    """images/s of the oracle port on the host cores for a bounded sample of the workload."""
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    t0 = time.time()
    probe = build_oracle_step(cfg_name, 2)
    probe()
    t1 = time.time()
    probe()
    t_img = (time.time() - t1) / 2
    n_steps = steps if steps is not None else 2
    b = int(max(1, min(32 if CONFIGS[cfg_name]["batch"] <= 32 else 16, budget_s / max(1e-6, (n_steps + warmup) * t_img))))
    step = build_oracle_step(cfg_name, b)
    for _ in range(warmup):
        step()
    ts = time.time()
    for _ in range(n_steps):
        step()
    dt = (time.time() - ts) / n_steps
    return dict(value=b / dt, unit="images/s", cores=torch.get_num_threads(), kind="port",
                sample=f"{CONFIGS[cfg_name]['desc']}, fp32, batch {b}, {n_steps} timed step(s) after {warmup} warm-up "
                       f"(oracle/peft_oracle.py on torch CPU; setup {t1 - t0:.0f}s)"), dt, b


def run_reference(args, rank):
    if rank != 0:
        return
    c = CONFIGS[args.config]
    steps = max(1, args.steps)
    base, dt, b = cpu_reference_rate(args.config, budget_s=150.0, steps=steps, warmup=max(1, min(args.warmup, 2)))
    line = {
        "impl": "reference",
        "metric": "fine-tune step images/s (fwd+bwd) ViT-B/16 KAdaptation" if args.config == "C3" else f"fine-tune step images/s (fwd+bwd) {c['desc']}",
        "value": base["value"], "unit": "images/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.config}: {c['desc']}, 224x224 synthetic, CPU sample batch {b}"},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", default="C3", choices=list(CONFIGS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override (default: global batch / N, or the global batch with --scaling weak)")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong (default): the config's GLOBAL batch is fixed and split over the ranks (SURVEY 8e); weak: per-GPU batch fixed")
    ap.add_argument("--depth", type=int, default=0, help="debug only: override depth (marks the line invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="fp32 = the verification mode (plain CUDA-core kernels behind the same engine); reported, not the headline")
    ap.add_argument("--fused-sgd", type=int, default=1, help="1 (default): one kernel does all-reduce + SGD-momentum + bucket zeroing "
                    "(peft_allreduce_sgd over symmetric memory); 0: NCCL all_reduce + torch.optim.SGD")
    ap.add_argument("--graph", type=int, default=1, help="1: replay the whole step from one CUDA graph (default), 0: eager launches")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(3, args.warmup)

    import torch.distributed as dist
    import torch.nn.functional as F
    from peft_b200 import synth                   # synthetic weights / inputs (nothing from oracle/ on this arm)
    from peft_b200 import _lib, ops
    from peft_b200.classifier import Classifier
    from peft_b200.dist import FusedSGDBucket, GradBucket

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # everything runs on a non-default stream: autograd binds a parameter's gradient-accumulation node to the stream of its
    # first use, and a node bound to the legacy default stream cannot take part in a later CUDA-graph capture
    torch.cuda.set_stream(torch.cuda.Stream(device=dev))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    c = dict(CONFIGS[args.config])
    depth = args.depth or c["depth"]
    if args.batch:
        Bpg = args.batch
    elif args.scaling == "strong":
        if c["batch"] % world:
            raise SystemExit(f"global batch {c['batch']} is not divisible by {world} ranks")
        Bpg = c["batch"] // world
    else:
        Bpg = c["batch"]
    cfg = vit_cfg(c, depth)
    sd = synth.make_state_dict(cfg, seed=0, grad_test=False)
    names = set(synth.trainable_names(cfg, sd, train_rule=c.get("phm_dim", 4) <= 64))   # few_shot freezes phm_rule ('W' filter)
    kw = adapter_kwargs(c)
    if c["dialect"] == "timm":
        from peft_b200.timm_vit import VisionTransformer
        vit = VisionTransformer(img_size=224, patch_size=c["patch"], num_classes=10, embed_dim=c["width"], depth=depth,
                                num_heads=c["heads"], qkv_bias=True, adapter=c["adapter"], **kw)
    else:
        from peft_b200.clip_model import VisionTransformer
        vit = VisionTransformer(224, c["patch"], c["width"], depth, c["heads"], c["out_dim"], adapter=c["adapter"], **kw)
    vit.load_state_dict(sd, strict=False)
    vit.precision = args.precision
    if c.get("fastfood_d"):
        from peft_b200.fastfood import FastfoodWrap
        import numpy as np
        torch.manual_seed(0)
        np.random.seed(0)
        for n_, p_ in vit.named_parameters():
            p_.requires_grad_(not n_.startswith("head."))
        vit = FastfoodWrap(vit.to(dev), c["fastfood_d"], device=local_rank)
        names = {"V"}
    model = Classifier(vit, cfg.feat_dim, 10, trainable_filter=lambda n: n in names).to(dev)
    head = synth.make_head(cfg)
    model.layers[0].load_state_dict({"weight": head["layers.0.weight"], "bias": head["layers.0.bias"]})
    model.train()
    from peft_b200.dist import kind_major
    trainable = kind_major([(n, p) for n, p in model.named_parameters() if p.requires_grad])   # per-layer tensors adjacent: see kind_major
    if args.fused_sgd:
        bucket = FusedSGDBucket(trainable, lr=1e-3, momentum=0.9, weight_decay=0.0)
        opt = None
    else:
        bucket = GradBucket(trainable)
        opt = torch.optim.SGD(trainable, lr=1e-3, momentum=0.9, weight_decay=0.0)

    g = torch.Generator().manual_seed(1234 + rank)
    n_host = 2
    host_imgs = [torch.randn(Bpg, 3, 224, 224, generator=g).pin_memory() for _ in range(n_host)]
    host_lbls = [torch.randint(0, 10, (Bpg,), generator=g).pin_memory() for _ in range(n_host)]
    dev_imgs = host_imgs[0].to(dev)
    dev_lbls = host_lbls[0].to(dev)

    def step(images, labels):
        bucket.zero()
        out = model(images)
        loss = F.cross_entropy(out, labels)
        loss.backward()
        if opt is None:
            bucket.step()
        else:
            bucket.all_reduce_mean()
            opt.step()
        return loss

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(dev_imgs, dev_lbls)
    sync_all()

    # ---- pass 1 (eager, instrumented): CUDA events around every tcgen05 GEMM launch of K steps -> roofline of the
    #      dominant kernel.  Eager launches are needed here because events cannot be read back from inside a graph.
    gemm_events = []
    orig_gemm = ops.gemm

    def timed_gemm(a, b, out, *a_, **k_):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        r = orig_gemm(a, b, out, *a_, **k_)
        e1.record()
        gemm_events.append((e0, e1, 2.0 * a.shape[0] * b.shape[0] * a.shape[1]))
        return r

    import peft_b200.engine as engine_mod
    engine_mod.ops.gemm = timed_gemm
    sync_all()
    p0 = torch.cuda.Event(enable_timing=True); p1 = torch.cuda.Event(enable_timing=True)
    launches0 = _lib.launch_count()
    p0.record()
    for _ in range(args.steps):
        loss = step(dev_imgs, dev_lbls)
    p1.record()
    sync_all()
    eager_launches = _lib.launch_count() - launches0
    engine_mod.ops.gemm = orig_gemm
    eager_ms = p0.elapsed_time(p1)
    gemm_ms = sum(e0.elapsed_time(e1) for e0, e1, _ in gemm_events)
    gemm_flops = sum(f for _, _, f in gemm_events)

    # ---- pass 2 (headline): the same step, replayed from ONE CUDA graph (peft_b200.graph) unless --graph 0
    graphed = None
    if args.graph:
        from peft_b200.graph import GraphedTrainStep
        del loss                        # drops the last eager autograd graph
        graphed = GraphedTrainStep(model, bucket, opt, dev_imgs, dev_lbls, warmup=2)

        def run_step(images, labels):
            return graphed(images, labels)
    else:
        run_step = step
    for _ in range(3):
        run_step(dev_imgs, dev_lbls)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    sync_all()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(args.steps):
        loss = run_step(None, None) if graphed is not None else step(dev_imgs, dev_lbls)
    t1.record()
    sync_all()
    launches = graphed.launches_per_step * args.steps if graphed is not None else eager_launches
    clocks = sampler.stop() if rank == 0 else None
    ms = t0.elapsed_time(t1)
    tms = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_step = float(tms.item()) / args.steps
    value = Bpg * world / (ms_step * 1e-3)

    # ---- end to end: host (pinned) images -> device every step (prefetched on a copy stream), loss read back
    e2e = None
    if not args.no_e2e:
        copy_stream = torch.cuda.Stream()
        bufs = [(torch.empty_like(dev_imgs), torch.empty_like(dev_lbls)) for _ in range(2)]
        ready = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]

        def prefetch(i):
            s = i % 2
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[s])
                bufs[s][0].copy_(host_imgs[i % n_host], non_blocking=True)
                bufs[s][1].copy_(host_lbls[i % n_host], non_blocking=True)
                ready[s].record(copy_stream)

        def e2e_loop(n):
            for s in range(2):
                consumed[s].record()
            prefetch(0)
            last = 0.0
            for i in range(n):
                if i + 1 < n:
                    prefetch(i + 1)
                s = i % 2
                torch.cuda.current_stream().wait_event(ready[s])
                l = run_step(bufs[s][0], bufs[s][1])
                consumed[s].record()
                last = float(l.item())          # the per-step host sync of train_one (kadaptation_clip.py:315)
            return last

        e2e_loop(2)
        sync_all()
        w0 = time.perf_counter()
        e2e_loop(args.steps)
        sync_all()
        w1 = time.perf_counter()
        te = torch.tensor([w1 - w0], device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": Bpg * world * args.steps / float(te.item()), "unit": "images/s",
               "h2d_bytes_per_step": int(dev_imgs.numel() * 4 + dev_lbls.numel() * 8) * world,
               "d2h_bytes_per_step": 4 * world}

    if rank == 0:
        pk = peaks()
        flop_img = c["gflop_img"] * 1e9 * (depth / c["depth"])
        line = {
            "metric": "fine-tune step images/s (fwd+bwd) ViT-B/16 KAdaptation" if args.config == "C3" else f"fine-tune step images/s (fwd+bwd) {c['desc']}",
            "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": f"{args.config}: {c['desc']}, 224x224 synthetic images, global batch {Bpg * world} = {world} x {Bpg} per GPU, SGD m=0.9, 10-class BN+Linear head",
                       "global_batch": Bpg * world, "seq_len": cfg.seq, "parallelism": f"dp{world}",
                       "l2": "per-step working set (GBs of saved activations) >> 126 MB L2; no flush needed",
                       "allreduce_bytes": bucket.nbytes,
                       "optimizer": "peft_allreduce_sgd (one-shot NVLink all-reduce + SGD m=0.9 + bucket zeroing, 1 kernel)" if opt is None
                       else "ncclAllReduce + torch.optim.SGD"},
            "step_frac_of_bf16_peak": value * flop_img / (pk["tflops_burst"] * 1e12 * world),
            "step_frac_of_bf16_peak_sustained": value * flop_img / (pk["tflops"] * 1e12 * world),
            "peaks": {"bf16_tflops_burst": pk["tflops_burst"], "bf16_tflops_sustained": pk["tflops"], "hbm_gbs": pk["hbm"],
                      "source": "MEASURED_PEAKS.json" if pk["src"] == "measured" else "fallback (B200_PROFILING.md)"},
            "launch_mode": "cuda-graph replay (1 graph launch per step)" if graphed is not None else "eager",
            "eager_ms_per_step": eager_ms / args.steps,
            "roofline": {"bound": "tensor", "kernel": "gemm_bf16_tcgen05_kernel (all dense products of the step)",
                         "achieved": gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else None, "peak": pk["tflops"],
                         "unit": "TFLOP/s", "frac": (gemm_flops / (gemm_ms * 1e-3) / 1e12) / pk["tflops"] if gemm_ms > 0 else None,
                         "traffic": ncu_gemm_traffic(args.config) if depth == c["depth"] and Bpg == c["batch"] else None,
                         "traffic_unit": "bytes per launch (mean over the 8 products of a layer, ncu dram read+write; profiles/gemm_traffic.json, null if the kernels changed since the capture)",
                         "frac_of_burst": (gemm_flops / (gemm_ms * 1e-3) / 1e12) / pk["tflops_burst"] if gemm_ms > 0 else None,
                         "peak_source": pk["src"] + " bf16_tflops_sustained",
                         "gemm_share_of_step": gemm_ms / eager_ms if eager_ms > 0 else None, "gemm_launches": len(gemm_events),
                         "measured_in": "eager instrumented pass of the same K steps (CUDA events per launch)"},
            "gpu_launches": int(launches), "clocks": clocks, "e2e": e2e,
            "final_loss": float(loss.item()),
        }
        if args.depth and args.depth != c["depth"]:
            line["invalid"] = "debug run with reduced depth"
        if args.precision == "fp32":
            line["dtype"] = "f32"
            line["note"] = "fp32 verification mode: plain CUDA-core kernels (csrc/f32_sm100.cuh), no tensor cores; not the headline path"
            line["roofline"]["kernel"] = "f32::gemm_kernel (CUDA-core SGEMM; fraction is against the bf16 tensor peak and only shows the gap)"
            line["roofline"]["traffic"] = None
        if world == 1 and not args.no_cpu_baseline:
            base, _, _ = cpu_reference_rate(args.config, budget_s=20.0, steps=2, warmup=1)
            line["cpu_baseline"] = base
            if args.config != "C1":
                # SURVEY 8d's stated CPU configuration, exactly: BASELINE.json configs[0] (CLIP ViT-B/32 + KAdaptation, batch 32)
                c1, _, _ = cpu_reference_rate("C1", budget_s=12.0, steps=2, warmup=1)
                line["cpu_baseline_c1"] = c1
        print(json.dumps(line), flush=True)
    if world > 1:
        # NCCL kernels captured into a live CUDA graph keep the communicator busy: destroy the graph first, then the
        # group.  A watchdog leaves the process if the teardown still blocks (every rank has its result by now).
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        wd = threading.Timer(20.0, lambda: os._exit(0))
        wd.daemon = True
        wd.start()
        if graphed is not None:
            graphed.release()
        del graphed
        torch.cuda.synchronize()
        dist.destroy_process_group()
        wd.cancel()


if __name__ == "__main__":
    main()
