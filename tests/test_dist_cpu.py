"""world_size-2 gloo test (CPU) of the flat gradient bucket: N-rank averaged gradients == gradients of the mean loss
over the concatenated batch."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from peft_b200.dist import GradBucket
    torch.manual_seed(0)
    lin = torch.nn.Linear(8, 3)
    frozen = torch.nn.Linear(8, 8)
    for p in frozen.parameters():
        p.requires_grad_(False)
    bucket = GradBucket(list(lin.parameters()) + list(frozen.parameters()))
    assert bucket.numel == 8 * 3 + 3
    x = torch.randn(6, 8, generator=torch.Generator().manual_seed(1))
    y = torch.randn(6, 3, generator=torch.Generator().manual_seed(2))
    xs, ys = x[rank * 3:(rank + 1) * 3], y[rank * 3:(rank + 1) * 3]
    for _ in range(2):  # second iteration checks zero() keeps the views attached
        bucket.zero()
        loss = ((lin(frozen(xs)) - ys) ** 2).mean()
        loss.backward()
        bucket.check_attached()
        bucket.all_reduce_mean()
    got = [p.grad.clone() for p in lin.parameters()]
    lin2 = torch.nn.Linear(8, 3)
    lin2.load_state_dict(lin.state_dict())
    ((lin2(frozen(x)) - y) ** 2).mean().backward()
    ok = all(torch.allclose(a, b.grad, atol=1e-6) for a, b in zip(got, lin2.parameters()))
    q.put((rank, ok))
    dist.destroy_process_group()


def test_bucket_allreduce_two_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res), res


def _bn_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from peft_b200.dist import GlobalBatchNorm1d
    x = torch.randn(8, 5, generator=torch.Generator().manual_seed(1))
    wgt = torch.randn(8, 5, generator=torch.Generator().manual_seed(2))
    xs = x[rank * 4:(rank + 1) * 4].clone().requires_grad_(True)
    bn = GlobalBatchNorm1d(5)
    y = bn(xs)
    (y * wgt[rank * 4:(rank + 1) * 4]).sum().backward()
    ref_bn = torch.nn.BatchNorm1d(5, affine=False)
    xr = x.clone().requires_grad_(True)
    yr = ref_bn(xr)
    (yr * wgt).sum().backward()
    ok = (torch.allclose(y, yr[rank * 4:(rank + 1) * 4], atol=1e-5) and torch.allclose(xs.grad, xr.grad[rank * 4:(rank + 1) * 4], atol=1e-5)
          and torch.allclose(bn.running_mean, ref_bn.running_mean, atol=1e-6) and torch.allclose(bn.running_var, ref_bn.running_var, atol=1e-5))
    q.put((rank, ok))
    dist.destroy_process_group()


def test_global_batchnorm_two_ranks_equals_single_rank():
    """with sync_bn the N-rank head equals the 1-rank head on the concatenated batch (values AND gradients)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_bn_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res), res


def _fused_worker(rank, world, port, q):
    """Host logic of FusedSGDBucket on CPU tensors (gloo): flat parameter / gradient / momentum views and the unfused
    arithmetic == torch.optim.SGD on the concatenated batch."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from peft_b200.dist import FusedSGDBucket
    torch.manual_seed(0)
    lin = torch.nn.Linear(8, 3)
    ref = torch.nn.Linear(8, 3)
    ref.load_state_dict(lin.state_dict())
    bucket = FusedSGDBucket(list(lin.parameters()), lr=0.1, momentum=0.9, weight_decay=1e-2)
    opt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-2)
    x = torch.randn(6, 8, generator=torch.Generator().manual_seed(1))
    y = torch.randn(6, 3, generator=torch.Generator().manual_seed(2))
    xs, ys = x[rank * 3:(rank + 1) * 3], y[rank * 3:(rank + 1) * 3]
    for _ in range(3):
        bucket.zero()
        ((lin(xs) - ys) ** 2).mean().backward()
        bucket.step()
        opt.zero_grad()
        ((ref(x) - y) ** 2).mean().backward()
        opt.step()
    ok = all(torch.allclose(a, b, atol=1e-6) for a, b in zip(lin.parameters(), ref.parameters()))
    ok = ok and float(bucket.flat.abs().max()) == 0.0
    q.put((rank, ok))
    dist.destroy_process_group()


def test_fused_sgd_bucket_two_ranks_cpu():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_fused_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res), res


def test_kind_major_order_and_stack_fallback():
    """dist.kind_major: the same tensor of consecutive layers becomes adjacent, kinds in order of first appearance, layer order kept;
    timm_vit.stack_params on CPU tensors (or any layout that is not stacked in memory) is plain torch.stack with ordinary
    per-parameter gradients."""
    import torch
    from peft_b200 import timm_vit
    from peft_b200.dist import kind_major
    names = ["vit.blocks.0.attn.W_left1", "vit.blocks.0.attn.W_right1", "vit.blocks.1.attn.W_left1", "vit.blocks.1.attn.W_right1",
             "vit.blocks.10.attn.W_left1", "vit.blocks.10.attn.W_right1", "layers.0.weight", "layers.0.bias", "vit.cls_token"]
    named = [(n, torch.nn.Parameter(torch.full((2, 3), float(i)))) for i, n in enumerate(names)]
    order = [names[[id(p) for _, p in named].index(id(q))] for q in kind_major(named)]
    assert order == ["vit.blocks.0.attn.W_left1", "vit.blocks.1.attn.W_left1", "vit.blocks.10.attn.W_left1",
                     "vit.blocks.0.attn.W_right1", "vit.blocks.1.attn.W_right1", "vit.blocks.10.attn.W_right1",
                     "layers.0.weight", "layers.0.bias", "vit.cls_token"]
    ps = [p for n, p in named if n.endswith("W_left1")]
    hits = timm_vit.STACK_FAST_HITS
    st = timm_vit.stack_params(ps)
    assert timm_vit.STACK_FAST_HITS == hits and torch.equal(st, torch.stack(ps))
    (st * torch.arange(3.0).reshape(3, 1, 1)).sum().backward()
    for i, p in enumerate(ps):
        assert torch.equal(p.grad, torch.full((2, 3), float(i)))
