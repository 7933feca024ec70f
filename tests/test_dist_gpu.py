"""SURVEY §8e correctness test on real GPUs (needs >= 2; skipped otherwise): one data-parallel KAdaptation step on N = 2
ranks over NCCL (per-rank half batch, GlobalBatchNorm head statistics, one flat all-reduce of the adapter + head
gradients) equals the 1-rank step on the concatenated batch.

fp32 mode: logits and averaged gradients <= 1e-5 relative.  bf16 mode: within the bf16 gate (2e-2) of the 1-rank result.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _build(precision, sync_bn, dev):
    from oracle import synth
    from oracle.synth import VitCfg
    from peft_b200.classifier import Classifier
    from peft_b200.timm_vit import VisionTransformer
    cfg = VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=64, patch=32, depth=2)
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    names = synth.trainable_names(cfg, sd)
    model = VisionTransformer(img_size=cfg.img, patch_size=cfg.patch, num_classes=cfg.num_classes, embed_dim=cfg.width,
                              depth=cfg.depth, num_heads=cfg.heads, qkv_bias=True, adapter="kron", phm_dim=4)
    model.load_state_dict(sd, strict=False)
    model.precision = precision
    clf = Classifier(model, cfg.feat_dim, cfg.num_classes, trainable_filter=lambda n: n in names, sync_bn=sync_bn).to(dev)
    head = synth.make_head(cfg)
    clf.layers[0].load_state_dict({"weight": head["layers.0.weight"], "bias": head["layers.0.bias"]})
    clf.train()
    return cfg, clf


def _step(clf, images, labels, bucket=None):
    if bucket is not None:
        bucket.zero()               # the gradients are views of the bucket's flat buffer: never reset them to None
    logits = clf(images)
    F.cross_entropy(logits, labels).backward()
    if bucket is not None:
        bucket.check_attached()
        bucket.all_reduce_mean()
    return logits.detach()


def _worker(rank, world, port, precision, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from oracle import synth
    from peft_b200.dist import GradBucket
    cfg, clf = _build(precision, True, dev)
    B = 16
    images, labels = synth.make_batch(cfg, B)
    lo, hi = rank * B // world, (rank + 1) * B // world
    trainable = [p for p in clf.parameters() if p.requires_grad]
    bucket = GradBucket(trainable)
    logits = _step(clf, images[lo:hi].to(dev), labels[lo:hi].to(dev), bucket)
    torch.cuda.synchronize()
    got = torch.cat([p.grad.detach().float().reshape(-1) for p in trainable]).cpu()
    out = None
    if rank == 0:
        # 1-rank step on the concatenated batch, same process, no collective.  The N-rank loss is the mean of per-rank
        # means over equal shards == the global mean, so averaged gradients must match.
        _, ref = _build(precision, False, dev)
        rl = _step(ref, images.to(dev), labels.to(dev))
        rt = [p for p in ref.parameters() if p.requires_grad]
        rg = torch.cat([p.grad.detach().float().reshape(-1) for p in rt]).cpu()
        rel = lambda a, b: float((a.double() - b.double()).norm() / (b.double().norm() + 1e-300))
        out = (rel(logits.cpu(), rl[lo:hi].cpu()), rel(got, rg))
    q.put((rank, out))
    dist.barrier()
    torch.cuda.synchronize()
    dist.destroy_process_group()


def _fused_worker(rank, world, port, q):
    """3 fused steps (peft_allreduce_sgd over symmetric memory) on 2 ranks vs 3 steps of torch.optim.SGD on the concatenated
    batch in one process: parameters after the steps must agree (fp32 mode), and be bit-identical across the ranks."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from oracle import synth
    from peft_b200.dist import FusedSGDBucket
    cfg, clf = _build("fp32", True, dev)
    B = 16
    images, labels = synth.make_batch(cfg, B)
    lo, hi = rank * B // world, (rank + 1) * B // world
    trainable = [p for p in clf.parameters() if p.requires_grad]
    bucket = FusedSGDBucket(trainable, lr=0.05, momentum=0.9, weight_decay=1e-3)
    for _ in range(3):
        bucket.zero()
        F.cross_entropy(clf(images[lo:hi].to(dev)), labels[lo:hi].to(dev)).backward()
        bucket.step()
    torch.cuda.synchronize()
    mine = torch.cat([p.detach().float().reshape(-1) for p in trainable])
    both = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(both, mine)
    out = None
    if rank == 0:
        _, ref = _build("fp32", False, dev)
        rt = [p for p in ref.parameters() if p.requires_grad]
        opt = torch.optim.SGD(rt, lr=0.05, momentum=0.9, weight_decay=1e-3)
        for _ in range(3):
            opt.zero_grad()
            F.cross_entropy(ref(images.to(dev)), labels.to(dev)).backward()
            opt.step()
        want = torch.cat([p.detach().float().reshape(-1) for p in rt])
        start = torch.cat([p.detach().float().reshape(-1) for p in [q_ for q_ in _build("fp32", False, dev)[1].parameters() if q_.requires_grad]])
        moved = float((want - start).norm())
        out = (float((mine - want).norm()) / (moved + 1e-30), bool(torch.equal(both[0], both[1])), moved)
    q.put((rank, out))
    dist.barrier()
    torch.cuda.synchronize()
    dist.destroy_process_group()


def test_two_rank_fused_allreduce_sgd():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_fused_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=120)
    err, identical, moved = res[0]
    print(f"[dist fused] 2-rank fused all-reduce + SGD vs 1-rank torch SGD after 3 steps: update error {err:.3e} of |update| {moved:.3e}, "
          f"ranks bit-identical: {identical}")
    assert identical and err < 1e-4, (err, identical)


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-5), ("bf16", 2e-2)])
def test_two_rank_step_equals_one_rank_step(precision, tol):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, precision, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=120)
    e_logits, e_grad = res[0]
    print(f"[dist {precision}] 2-rank vs 1-rank: logits rel {e_logits:.3e}, averaged gradient rel {e_grad:.3e}")
    assert e_logits < tol and e_grad < tol, (e_logits, e_grad)
