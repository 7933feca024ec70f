"""Timing helper (not a test): peft_colsum_bf16 at the C5 shapes (12 800 tokens)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from peft_b200 import ops
from bench_ops import timeit, graph_time
for N in (768, 2304, 3072):
    x = torch.randn(12800, N, device="cuda").to(torch.bfloat16)
    out = torch.zeros(N, device="cuda")
    t = graph_time(lambda: ops.colsum(x, out))
    print(f"colsum [12800, {N}] PEFT_COLSUM_CTAS={os.environ.get('PEFT_COLSUM_CTAS', '2')}: {t:6.1f} us  {12800 * N * 2 / t / 1e3:7.1f} GB/s", flush=True)
