"""GPU input pipeline: peft_preprocess_clip must be BIT-IDENTICAL to the reference's torchvision-on-PIL transform
(FS/clip_load.py:80-87; committed goldens) — byte/integer work, so the bar is exact equality, not a tolerance."""
import pytest
import torch

from oracle import synth
from oracle.preprocess_oracle import clip_transform

pytestmark = pytest.mark.gpu

PRE_CASES = ["preprocess_300x400_96", "preprocess_500x375_128", "preprocess_150x200_224", "preprocess_97x97_64", "preprocess_640x481_96"]


@pytest.mark.parametrize("name", PRE_CASES)
def test_preprocess_matches_reference_golden(golden, name):
    from peft_b200.preprocess import ClipPreprocess
    c = golden[name]
    pre = ClipPreprocess(c["n_px"])
    out = pre(synth.make_image_u8(c["h"], c["w"]).cuda())
    assert torch.equal(out.cpu(), c["out"])


@pytest.mark.parametrize("h,w,n_px", [(1080, 1920, 224), (224, 224, 224), (231, 1001, 224), (2000, 300, 336)])
def test_preprocess_matches_oracle_other_sizes(h, w, n_px):
    """Large downscales (long coefficient rows), identity size, extreme aspect ratios; and a batch of mixed sizes."""
    from peft_b200.preprocess import ClipPreprocess
    img = synth.make_image_u8(h, w, seed=7)
    pre = ClipPreprocess(n_px)
    assert torch.equal(pre(img.cuda()).cpu(), clip_transform(img.numpy(), n_px))


def test_preprocess_batch_and_no_cpu_fallback():
    from peft_b200.preprocess import ClipPreprocess
    pre = ClipPreprocess(96)
    imgs = [synth.make_image_u8(h, w, seed=3) for h, w in ((120, 160), (300, 200), (120, 160))]
    out = pre.batch([i.cuda() for i in imgs])
    for i, im in enumerate(imgs):
        assert torch.equal(out[i].cpu(), clip_transform(im.numpy(), 96))
    with pytest.raises(RuntimeError):
        pre(imgs[0])
