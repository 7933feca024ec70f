"""Oracle of the l2 search — TEST INFRASTRUCTURE.  Line-by-line restatement of `hyperparameter_sweep`
(few_shot/vision_benchmark/evaluation/kadaptation_clip.py:171-223) with the training run replaced by a callable; the trainer
module itself cannot be imported here (yacs / vision_datasets / nltk), so parity is pinned on the control flow:
same candidates, same evaluation order (including the committed `[left]` quirk at :207), same returned peak."""
import numpy as np


def hyperparameter_sweep(train_task, lower, upper):
    l2_lambda_list = np.logspace(lower, upper, num=97).tolist()
    l2_lambda_init_idx = [i for i, val in enumerate(l2_lambda_list) if val in set(np.logspace(lower, upper, num=7))]
    peak_idx = -1
    peak_score = 0
    evaluated = []
    for idx in l2_lambda_init_idx:
        wd = l2_lambda_list[idx]
        try:
            evaluated.append(wd)
            best_score_ = train_task(wd)
        except Exception:
            best_score_ = 0.0
            continue
        if best_score_ > peak_score:
            peak_idx = idx
            peak_score = best_score_
    step_span = 8
    while step_span > 0:
        left, right = max(peak_idx - step_span, 0), min(peak_idx + step_span, len(l2_lambda_list) - 1)
        search_idx = []
        if left != peak_idx:
            search_idx.append(left)
        if right != peak_idx:
            search_idx.append(right)
        for idx in search_idx:
            wd = l2_lambda_list[left]
            try:
                evaluated.append(wd)
                best_score_ = train_task(wd)
            except Exception:
                best_score_ = 0.0
                continue
            if best_score_ > peak_score:
                peak_idx = idx
                peak_score = best_score_
        step_span //= 2
    return l2_lambda_list[peak_idx], peak_score, evaluated
