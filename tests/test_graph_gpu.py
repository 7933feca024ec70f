"""Whole-step CUDA-graph capture: a replayed step must equal the eager step (same kernels, same order)."""
import copy

import pytest
import torch
import torch.nn.functional as F

from oracle import synth
from oracle.synth import VitCfg

pytestmark = pytest.mark.gpu


def make(cfg):
    from peft_b200.classifier import Classifier
    from peft_b200.dist import GradBucket
    from peft_b200.timm_vit import VisionTransformer
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    names = set(synth.trainable_names(cfg, sd))
    vit = VisionTransformer(img_size=cfg.img, patch_size=cfg.patch, num_classes=10, embed_dim=cfg.width, depth=cfg.depth,
                            num_heads=cfg.heads, qkv_bias=True, adapter=cfg.adapter, phm_dim=4, phm_rank=1)
    vit.load_state_dict(sd, strict=False)
    model = Classifier(vit, cfg.feat_dim, 10, trainable_filter=lambda n: n in names).cuda()
    head = synth.make_head(cfg)
    model.layers[0].load_state_dict({"weight": head["layers.0.weight"], "bias": head["layers.0.bias"]})
    model.train()
    trainable = [p for p in model.parameters() if p.requires_grad]
    bucket = GradBucket(trainable)
    opt = torch.optim.SGD(trainable, lr=1e-2, momentum=0.9)
    return model, bucket, opt, trainable


def test_graph_replay_matches_eager():
    from peft_b200.graph import GraphedTrainStep
    cfg = VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=64, patch=32, depth=2)
    B = 16
    batches = [synth.make_batch(cfg, B, seed_img=100 + i, seed_lbl=200 + i) for i in range(3)]
    # eager: 3 steps
    model, bucket, opt, params = make(cfg)
    eager_losses = []
    for im, lb in batches:
        bucket.zero()
        loss = F.cross_entropy(model(im.cuda()), lb.cuda())
        loss.backward()
        opt.step()
        eager_losses.append(float(loss))
    eager_params = [p.detach().clone() for p in params]
    # graphed: warm-up steps run on a scratch copy of the state, then the same 3 batches are replayed
    model, bucket, opt, params = make(cfg)
    init = [p.detach().clone() for p in params]
    bn = copy.deepcopy(model.channel_bn.state_dict())
    g = GraphedTrainStep(model, bucket, opt, batches[0][0].cuda(), batches[0][1].cuda(), warmup=2)
    assert g.launches_per_step > 20
    with torch.no_grad():                      # undo the warm-up / capture steps: parameters, momentum, BN statistics
        for p, p0 in zip(params, init):
            p.copy_(p0)
        for st in opt.state.values():
            st["momentum_buffer"].zero_()
        model.channel_bn.load_state_dict(bn)
    losses = [float(g(im.cuda(), lb.cuda())) for im, lb in batches]
    for a, b in zip(losses, eager_losses):
        assert abs(a - b) < 2e-3 * max(1.0, abs(b)), (losses, eager_losses)
    for p, q in zip(params, eager_params):
        assert float((p - q).norm() / (q.norm() + 1e-12)) < 2e-3
