"""CPU tests of the SURVEY §8f-4 host logic: the input-pipeline oracle against the real torchvision-on-PIL outputs (bit-equal),
the product's resampling-coefficient tables against the oracle's, and the l2 sweep driver against its oracle."""
import math

import numpy as np
import pytest
import torch

from oracle import synth
from oracle.preprocess_oracle import _coeffs, clip_transform
from oracle.sweep_oracle import hyperparameter_sweep

PRE_CASES = ["preprocess_300x400_96", "preprocess_500x375_128", "preprocess_150x200_224", "preprocess_97x97_64", "preprocess_640x481_96"]


@pytest.mark.parametrize("name", PRE_CASES)
def test_preprocess_oracle_is_bit_identical_to_torchvision_on_pil(golden, name):
    c = golden[name]
    out = clip_transform(synth.make_image_u8(c["h"], c["w"]).numpy(), c["n_px"])
    assert torch.equal(out, c["out"])


@pytest.mark.parametrize("n_in,n_out", [(400, 128), (375, 96), (200, 298), (97, 64), (224, 224), (33, 224)])
def test_product_coefficient_tables_match_oracle(n_in, n_out):
    from peft_b200.preprocess import resample_coeffs
    bounds, kk = resample_coeffs(n_in, n_out)
    ref = _coeffs(n_in, n_out)
    assert bounds.shape == (n_out, 2)
    for i, (xmin, k) in enumerate(ref):
        assert bounds[i, 0] == xmin and bounds[i, 1] == len(k)
        assert np.array_equal(kk[i, :len(k)].astype(np.int64), k)
        assert not kk[i, len(k):].any()


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
@pytest.mark.parametrize("faithful", [True, False])
def test_l2_search_follows_the_reference_control_flow(seed, faithful):
    from peft_b200.sweep import l2_search
    rng = np.random.RandomState(seed)
    centre, width = rng.uniform(-4, 4), rng.uniform(0.5, 3)
    calls = []

    def score(l2):                       # smooth unimodal score in log10(l2) + a trial that "crashes"
        calls.append(l2)
        if seed == 3 and len(calls) == 2:
            raise RuntimeError("simulated failed trial")
        return 50.0 + 40.0 * math.exp(-((math.log10(l2) - centre) / width) ** 2)

    best, peak, log = l2_search(score, -6, 6, faithful=faithful)
    if faithful:
        calls2 = []

        def score2(l2):
            calls2.append(l2)
            if seed == 3 and len(calls2) == 2:
                raise RuntimeError("simulated failed trial")
            return 50.0 + 40.0 * math.exp(-((math.log10(l2) - centre) / width) ** 2)

        rbest, rpeak, evaluated = hyperparameter_sweep(score2, -6, 6)
        assert [l for l, _ in log] == evaluated
        assert best == rbest and peak == rpeak
    else:
        assert abs(math.log10(best) - centre) < 12 / 96 + 1e-9 or peak >= 89.9      # converges onto the grid point next to the optimum
    assert any(isinstance(s, Exception) for _, s in log) == (seed == 3)


def test_shared_backbone_trials_reset_state():
    from peft_b200.sweep import SharedBackboneTrials
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    w0 = net[0].weight.detach().clone()
    ctx = SharedBackboneTrials(net)

    def train_and_score(model, l2):
        opt = torch.optim.SGD(model.parameters(), lr=0.1, weight_decay=l2)
        for _ in range(3):
            opt.zero_grad()
            model(torch.ones(5, 4)).sum().backward()
            opt.step()
        return float(model[0].weight.sum())

    a = ctx.run(train_and_score, [0.0, 0.1, 0.0])
    assert a[0][1] == a[2][1] != a[1][1]            # a trial starts from the same state whatever ran before it
    ctx.reset()
    assert torch.equal(net[0].weight, w0) and float(net[1].running_mean.abs().max()) == 0.0
