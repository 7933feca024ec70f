"""Fastfood projection on the GPU against the oracle restatement of FL/intrinsic/fastfood.py.

Tolerance: the path is fp32 end to end (north_star fp32 gate 1e-5 relative); the FWHT sums 2^22 terms in a different
order than the reference's chunk/cat recursion, so the bound used is 2e-5 relative (Frobenius) per tensor.
"""
import numpy as np
import pytest
import torch

from oracle import peft_oracle as O

pytestmark = pytest.mark.gpu
TOL = 2e-5


def rel(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def make_plan(shapes, d, seed=0):
    from peft_b200.fastfood import FastfoodPlan
    gn = np.random.RandomState(seed)
    gt = torch.Generator().manual_seed(seed)
    params = [O.fastfood_vars(int(np.prod(s)) if len(s) else 1, gn, gt) for s in shapes]
    plan = FastfoodPlan(shapes, params, d, "cuda")
    theta0 = [torch.randn(s, generator=gt) for s in shapes]
    plan.set_theta0(theta0)
    return plan, params, theta0


CASES = {
    "tiny_d16": ([(), (10,), (16,), (3, 7), (1000,)], 16),
    "small_d1000": ([(768,), (1000,), (1024,), (50, 768), (5000,), (8192,), (8193,)], 1000),
    "odd_d100": ([(20000,), (3, 4, 5), (70000,)], 100),
    "mid_d1000": ([(768, 768), (768,), (512, 768)], 1000),
    "big_d1000": ([(3072, 768), (2304, 768), (768, 3, 32, 32)], 1000),
    "d4096": ([(100000,), (5000,)], 4096),
}


@pytest.mark.parametrize("name", list(CASES))
def test_fastfood_forward_backward(name):
    shapes, d = CASES[name]
    plan, params, theta0 = make_plan(shapes, d)
    g = torch.Generator().manual_seed(5)
    V = torch.randn(d, generator=g) * 0.01
    saved0 = plan.theta0.clone()
    plan.theta0.zero_()                      # the projected update alone (no cancellation against theta0 ~ 1)
    rays = [t.clone() for t in plan.views(plan.project(V.cuda()))]
    plan.theta0.copy_(saved0)
    theta = plan.project(V.cuda())
    got = [t.clone() for t in plan.views(theta)]
    cots = [torch.randn(s, generator=g) for s in shapes]
    for dst, c in zip(plan.views(plan.dtheta), cots):
        dst.copy_(c)
    dV = plan.project_backward(plan.dtheta)
    torch.cuda.synchronize()
    Vr = V.clone().requires_grad_(True)
    loss = 0
    for s, p, t0, c, gt, ry in zip(shapes, params, theta0, cots, got, rays):
        DD = int(np.prod(s)) if len(s) else 1
        ray = O.fastfood_torched(Vr, DD, p).view(s)
        assert rel(ry, ray) < TOL, (s, rel(ry, ray))
        assert rel(gt, t0 + ray) < 1e-6
        loss = loss + (ray * c).sum()
    loss.backward()
    assert rel(dV, Vr.grad) < TOL, rel(dV, Vr.grad)


def test_fastfood_linearity_full_vit_b32():
    """Full ViT-B/32 visual parameter set (152 tensors, sum DD = 87.8 M): size-independent properties.
    (a) linearity: P(aV1 + bV2) - theta0 == a (P(V1) - theta0) + b (P(V2) - theta0);
    (b) adjointness: <P(V) - theta0, C> == <V, P^T(C)> for a random cotangent C."""
    from peft_b200.fastfood import FastfoodPlan
    D = 768
    shapes = [(D, 3, 32, 32), (D,), (50, D), (D,), (D,)]
    for _ in range(12):
        shapes += [(D,), (D,), (3 * D, D), (3 * D,), (D, D), (D,), (D,), (D,), (4 * D, D), (4 * D,), (D, 4 * D), (D,)]
    shapes += [(D,), (D,), (D, 512)]
    gn = np.random.RandomState(0)
    gt = torch.Generator().manual_seed(0)
    params = []
    for s in shapes:
        DD = int(np.prod(s))
        ll = int(np.ceil(np.log(DD) / np.log(2)))
        LL = 2 ** ll
        BB = (torch.rand(LL, generator=gt) * 2).long().float() * 2 - 1
        Pi = torch.from_numpy(gn.permutation(LL)).long()
        GG = torch.randn(LL, generator=gt)
        params.append([BB, Pi, GG, torch.sqrt(LL * torch.sum(GG ** 2)), LL])
    d = 1000
    plan = FastfoodPlan(shapes, params, d, "cuda")
    assert plan.T == 152
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(1)
    V1 = torch.randn(d, device=dev, generator=g)
    V2 = torch.randn(d, device=dev, generator=g)
    r1 = plan.project(V1).clone()
    r2 = plan.project(V2).clone()
    r3 = plan.project(0.5 * V1 - 2.0 * V2).clone()
    assert rel(r3, 0.5 * r1 - 2.0 * r2) < TOL
    C = torch.randn(plan.total_dd, device=dev, generator=g)
    for off, n in zip(plan.offsets, plan.numels):      # padding gaps between tensors carry no gradient
        pass
    mask = torch.zeros(plan.total_dd, device=dev)
    for off, n in zip(plan.offsets, plan.numels):
        mask[off:off + n] = 1
    C = C * mask
    plan.dtheta.copy_(C)
    dV = plan.project_backward(plan.dtheta)
    lhs = float((r1.double() * C.double()).sum())
    rhs = float((V1.double() * dV.double()).sum())
    assert abs(lhs - rhs) <= 1e-4 * max(abs(lhs), abs(rhs), 1e-3), (lhs, rhs)


def test_fastfood_wrap_module():
    """FastfoodWrap end to end on a small torch module: logits and V.grad against the oracle projection + autograd."""
    import torch.nn as nn
    from peft_b200.fastfood import FastfoodWrap
    torch.manual_seed(0)
    np.random.seed(0)
    net = nn.Sequential(nn.Linear(40, 300), nn.Tanh(), nn.Linear(300, 10))
    ref_net = nn.Sequential(nn.Linear(40, 300), nn.Tanh(), nn.Linear(300, 10))
    ref_net.load_state_dict(net.state_dict())
    wrap = FastfoodWrap(net.cuda(), 64, device=0)
    with torch.no_grad():
        wrap.V.copy_(torch.randn(64, generator=torch.Generator().manual_seed(3)) * 0.1)
    x = torch.randn(16, 40, generator=torch.Generator().manual_seed(4))
    out = wrap(x.cuda())
    out.square().sum().backward()
    # oracle: same B/Pi/G (taken from the wrapper), reference formula
    Vr = wrap.V.detach().cpu().clone().requires_grad_(True)
    sd = {}
    for name, p in ref_net.named_parameters():
        BB, Pi, GG, divisor, LL = wrap.fastfood_params[name]
        sd[name] = p.detach() + O.fastfood_torched(Vr, p.numel(), [BB, Pi, GG, divisor, LL]).view(p.shape)
    h = torch.tanh(x @ sd["0.weight"].T + sd["0.bias"])
    ref = h @ sd["2.weight"].T + sd["2.bias"]
    ref.square().sum().backward()
    assert rel(out, ref) < 1e-4
    assert rel(wrap.V.grad, Vr.grad) < 1e-4
    assert [n for n, _ in wrap.named_parameters()] == ["V"]
