"""Hyper-parameter sweep driver that shares ONE frozen backbone between trials.

replaces: `hyperparameter_sweep` (few_shot/vision_benchmark/evaluation/kadaptation_clip.py:171-223: the l2 search over 97
log-spaced values — 7 coarse points, then refinement with spans 8, 4, 2, 1 around the running peak) and the per-trial model
re-build of `train_task` (:226-278, a fresh `Classifier` = a fresh backbone load per trial, ~90 times per dataset).  Here a
trial resets the trainable tensors (adapter factors, head, optimiser state) of one live model to their initial values: the
engine, its packed frozen weights, workspaces, TMA descriptors and the captured step graph survive from trial to trial.

Host logic only (no arithmetic of its own): the training step it drives is the CUDA engine's.
"""
from __future__ import annotations

import numpy as np
import torch


def l2_candidates(log_lower: float, log_upper: float):
    """(the 97 candidates, indices of the 7 coarse starting points) — kadaptation_clip.py:174-176."""
    grid = np.logspace(log_lower, log_upper, num=97).tolist()
    coarse = set(np.logspace(log_lower, log_upper, num=7))
    return grid, [i for i, v in enumerate(grid) if v in coarse]


def l2_search(run_trial, log_lower: float = -6.0, log_upper: float = 6.0, faithful: bool = True):
    """The search of kadaptation_clip.py:171-223.  run_trial(l2) -> score (higher is better); a trial that raises scores 0
    (the reference's bare `except`, :185-189 — here the exception is recorded in the returned log instead of vanishing).
    faithful=True keeps the committed refinement quirk (:207 evaluates `l2_lambda_list[left]` for BOTH neighbours);
    faithful=False evaluates the neighbour it indexes.  Returns (best l2, best score, [(l2, score | exception)])."""
    grid, coarse = l2_candidates(log_lower, log_upper)
    log = []

    def trial(l2):
        try:
            s = float(run_trial(l2))
            log.append((l2, s))
            return s
        except Exception as exc:           # noqa: BLE001 — mirrors the reference's control flow, but keeps the evidence
            log.append((l2, exc))
            return None

    peak_idx, peak_score = -1, 0.0
    for idx in coarse:
        s = trial(grid[idx])
        if s is not None and s > peak_score:
            peak_idx, peak_score = idx, s
    span = 8
    while span > 0:
        left, right = max(peak_idx - span, 0), min(peak_idx + span, len(grid) - 1)
        for idx in [i for i in (left, right) if i != peak_idx]:
            s = trial(grid[left] if faithful else grid[idx])
            if s is not None and s > peak_score:
                peak_idx, peak_score = idx, s
        span //= 2
    return grid[peak_idx], peak_score, log


class SharedBackboneTrials:
    """Re-usable trial context: snapshot of every trainable tensor of `model` (adapter factors + head) taken once; `reset()`
    restores it in place — parameter storage, gradient-bucket views and captured CUDA graphs stay valid — so a new trial costs
    a few small copies instead of a backbone reload."""

    def __init__(self, model: torch.nn.Module, bucket=None):
        self.model, self.bucket = model, bucket
        self._init = [(p, p.detach().clone()) for p in model.parameters() if p.requires_grad]
        self._buffers = [(b, b.detach().clone()) for b in model.buffers()]       # BatchNorm running statistics of the head

    @torch.no_grad()
    def reset(self):
        for p, v in self._init:
            p.copy_(v)
        for b, v in self._buffers:
            b.copy_(v)
        if self.bucket is not None:
            self.bucket.flat.zero_()
            if hasattr(self.bucket, "flat_mom"):
                self.bucket.flat_mom.zero_()

    def run(self, train_and_score, l2_values):
        """[(l2, score)] — `train_and_score(model, l2)` trains the (reset) model and returns its validation score."""
        out = []
        for l2 in l2_values:
            self.reset()
            out.append((l2, float(train_and_score(self.model, l2))))
        return out
