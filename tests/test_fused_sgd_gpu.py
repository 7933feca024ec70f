"""peft_allreduce_sgd at world size 1 (the 2-rank exchange is tests/test_dist_gpu.py): the fused kernel must reproduce
torch.optim.SGD(momentum, dampening 0, weight_decay, nesterov False) — the optimiser the reference builds at
few_shot/vision_benchmark/optim/build.py:120-159 — over several steps, leave the bucket zeroed, keep every parameter a view of
the flat buffers, follow LR changes without re-capture, and survive CUDA-graph replay."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _params(seed=0):
    g = torch.Generator().manual_seed(seed)
    shapes = [(4, 4, 4), (4, 192, 1), (4, 1, 192), (10, 768), (10,), (7,)]      # total not a multiple of 4: exercises the padding
    return [torch.nn.Parameter(torch.randn(s, generator=g).cuda()) for s in shapes]


@pytest.mark.parametrize("mu,wd", [(0.9, 0.0), (0.9, 1e-2), (0.0, 0.0)])
def test_fused_sgd_matches_torch_sgd(mu, wd):
    from peft_b200.dist import FusedSGDBucket
    ours, ref = _params(), _params()
    bucket = FusedSGDBucket(ours, lr=0.05, momentum=mu, weight_decay=wd)
    opt = torch.optim.SGD(ref, lr=0.05, momentum=mu, weight_decay=wd)
    g = torch.Generator().manual_seed(5)
    for step in range(6):
        if step == 3:
            bucket.set_lr(0.01)
            for grp in opt.param_groups:
                grp["lr"] = 0.01
        bucket.zero()
        bucket.check_attached()
        for p, r in zip(ours, ref):
            gr = torch.randn(p.shape, generator=g).cuda()
            p.grad.add_(gr)                    # backward accumulates into the bucket views
            r.grad = gr.clone()
        bucket.step()
        opt.step()
        assert float(bucket.flat.abs().max()) == 0.0          # left zeroed for the next backward
        for p, r in zip(ours, ref):
            assert torch.allclose(p, r, rtol=1e-6, atol=1e-7), step
    base = bucket.flat_param.untyped_storage().data_ptr()
    assert all(p.data.untyped_storage().data_ptr() == base for p in ours)


def test_fused_sgd_in_cuda_graph():
    from peft_b200.dist import FusedSGDBucket
    ours, ref = _params(1), _params(1)
    torch.cuda.set_stream(torch.cuda.Stream())
    bucket = FusedSGDBucket(ours, lr=0.1, momentum=0.9)
    opt = torch.optim.SGD(ref, lr=0.1, momentum=0.9)
    gstat = [torch.randn_like(p) for p in ours]

    def step():
        for p, gr in zip(ours, gstat):
            p.grad.add_(gr)
        bucket.step()

    step()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    for _ in range(4):                          # 1 eager step + 3 replays (the capture itself executes nothing)
        for r, gr in zip(ref, gstat):
            r.grad = gr.clone()
        opt.step()
    for p, r in zip(ours, ref):
        assert torch.allclose(p, r, rtol=1e-5, atol=1e-6)
    torch.cuda.set_stream(torch.cuda.default_stream())


@pytest.mark.parametrize("adapter", ["kron", "lora"])
def test_stacked_parameter_views_match_per_parameter_accumulation(adapter):
    """dist.kind_major + timm_vit.stack_params: with the per-layer adapter tensors adjacent in the FusedSGDBucket the tower takes
    them as zero-copy stacked views and adds their stacked gradient to the bucket with one kernel per tensor kind; the bucket
    must then hold exactly what torch.stack + one AccumulateGrad per parameter produce, and three training steps must agree."""
    import torch.nn.functional as F
    from oracle import synth
    from oracle.synth import VitCfg
    from peft_b200 import timm_vit
    from peft_b200.classifier import Classifier
    from peft_b200.dist import FusedSGDBucket, kind_major
    cfg = VitCfg(dialect="timm", adapter=adapter, phm_dim=4, img=64, patch=32, depth=3)

    def build(ordered):
        sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
        names = set(synth.trainable_names(cfg, sd))
        vit = timm_vit.VisionTransformer(img_size=cfg.img, patch_size=cfg.patch, num_classes=10, embed_dim=cfg.width, depth=cfg.depth,
                                         num_heads=cfg.heads, qkv_bias=True, adapter=cfg.adapter, phm_dim=4, phm_rank=1)
        vit.load_state_dict(sd, strict=False)
        model = Classifier(vit, cfg.feat_dim, 10, trainable_filter=lambda n: n in names).cuda()
        head = synth.make_head(cfg)
        model.layers[0].load_state_dict({"weight": head["layers.0.weight"], "bias": head["layers.0.bias"]})
        model.train()
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        params = kind_major(named) if ordered else [p for _, p in named]
        return model, FusedSGDBucket(params, lr=0.05, momentum=0.9), dict(named)

    batches = [synth.make_batch(cfg, 8, seed_img=300 + i, seed_lbl=400 + i) for i in range(3)]
    ref_model, ref_bucket, ref_named = build(False)
    fast_model, fast_bucket, fast_named = build(True)
    for step, (im, lb) in enumerate(batches):
        h0 = timm_vit.STACK_FAST_HITS
        F.cross_entropy(ref_model(im.cuda()), lb.cuda()).backward()
        assert timm_vit.STACK_FAST_HITS == h0                      # layer-major bucket: plain torch.stack
        F.cross_entropy(fast_model(im.cuda()), lb.cuda()).backward()
        assert timm_vit.STACK_FAST_HITS - h0 == (5 if adapter == "kron" else 4)      # q / v x down / up factors (+ the trained phm_rule)
        for n, p in ref_named.items():
            assert torch.equal(p.grad, fast_named[n].grad), (step, n)
        ref_bucket.step()
        fast_bucket.step()
        for n, p in ref_named.items():
            assert torch.equal(p, fast_named[n]), (step, n)
