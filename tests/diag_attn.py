"""Diagnosis helper (not a test): per-head error map of the attention backward against fp32 torch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from peft_b200 import _lib, ops

def run(B, H, L, algo):
    _lib.check(_lib.load().peft_set_attention_algo(algo))
    D = H * 64
    g = torch.Generator(device="cuda").manual_seed(1)
    qkv = torch.randn(B * L, 3 * D, device="cuda", generator=g).to(torch.bfloat16)
    d_o = (torch.randn(B * L, D, device="cuda", generator=g) * 0.1).to(torch.bfloat16)
    o = torch.empty(B * L, D, device="cuda", dtype=torch.bfloat16)
    lse2 = torch.empty(B * H * L, device="cuda")
    delta = torch.empty(B * H * L, device="cuda")
    dqkv = torch.zeros(B * L, 3 * D, device="cuda", dtype=torch.bfloat16)
    scale = 0.125
    ops.attention_fwd(qkv, o, lse2, B, H, L, scale)
    ops.attention_bwd(qkv, o, d_o, lse2, delta, dqkv, B, H, L, scale)
    torch.cuda.synchronize()
    t = qkv.float().reshape(B, L, 3, H, 64).permute(2, 0, 3, 1, 4).clone().requires_grad_(True)
    s = (t[0] @ t[1].transpose(-2, -1)) * scale
    ref = (F.softmax(s, dim=-1) @ t[2]).transpose(1, 2).reshape(B * L, D)
    ref.backward(d_o.float())
    rd = t.grad                                                   # [3, B, H, L, 64]
    got = dqkv.float().reshape(B, L, 3, H, 64).permute(2, 0, 3, 1, 4)
    err = ((got - rd).flatten(3).norm(dim=3) / (rd.flatten(3).norm(dim=3) + 1e-20)).reshape(3, B * H)   # per (tensor, head)
    for i, nm in enumerate("qkv"):
        bad = (err[i] > 0.03).nonzero().flatten().tolist()
        print(f"B={B} H={H} L={L} algo={algo} d{nm}: max head error {float(err[i].max()):.4f}, bad heads {len(bad)} of {B * H}: {bad[:40]}", flush=True)

if __name__ == "__main__":
    for (B, H, L) in ((20, 12, 50), (3, 12, 50), (40, 12, 50), (26, 12, 197)):
        run(B, H, L, 3)
