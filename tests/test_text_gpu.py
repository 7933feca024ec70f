"""CLIP text tower + zero-shot head initialisation (FS/model.py:1303-1327, FS/feature.py:351-512,
FS/kadaptation_clip.py:128-138) through the fp32 C-ABI kernels (causal attention = peft_attention_fwd_f32_ex), against the
committed outputs of the unmodified reference CLIP.encode_text and against the oracle; tolerance 1e-5 (fp32 path)."""
import pytest
import torch

from oracle import peft_oracle as O
from oracle import synth

pytestmark = pytest.mark.gpu


def rel(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-300))


def _clip(cfg):
    from peft_b200.clip_model import CLIP
    m = CLIP(embed_dim=cfg["embed_dim"], image_resolution=64, vision_layers=1, vision_width=768, vision_patch_size=32,
             context_length=cfg["context"], vocab_size=cfg["vocab"], transformer_width=cfg["width"],
             transformer_heads=cfg["width"] // 64, transformer_layers=cfg["layers"])
    missing, unexpected = m.load_state_dict(synth.make_text_state_dict(**cfg), strict=False)
    assert not unexpected
    return m.cuda().eval()


def test_encode_text_matches_reference_golden(golden):
    c = golden["clip_text"]
    model = _clip(c["cfg"])
    toks = synth.make_tokens(c["n_tokens"], c["cfg"]["vocab"], c["cfg"]["context"])
    feats = model.encode_text(toks.cuda())
    assert rel(feats, c["features"]) < 1e-5, rel(feats, c["features"])
    z = model.zeroshot_head_weights([toks[4 * i:4 * i + 4].cuda() for i in range(3)])
    assert rel(z.t(), c["zeroshot"]) < 1e-5


def test_causal_mask_really_masks():
    """Tokens AFTER the EOT position must not influence the embedding (causal attention + EOT-row readout)."""
    cfg = dict(width=512, layers=2, vocab=1000, context=77, embed_dim=512)
    model = _clip(cfg)
    toks = synth.make_tokens(6, 1000, 77)
    a = model.encode_text(toks.cuda())
    t2 = toks.clone()
    for i in range(6):
        e = int(t2[i].argmax())
        t2[i, e + 1:] = torch.randint(1, 500, (77 - e - 1,))          # garbage after EOT (smaller ids: argmax unchanged)
    b = model.encode_text(t2.cuda())
    assert rel(a, b) < 1e-6
    t3 = toks.clone()
    t3[:, 1] = (t3[:, 1] + 7) % 900 + 1                                # a token BEFORE the EOT does matter
    assert rel(model.encode_text(t3.cuda()), a) > 1e-3


@pytest.mark.parametrize("width,layers,n", [(768, 1, 5), (512, 3, 130)])
def test_encode_text_matches_oracle_other_geometries(width, layers, n):
    cfg = dict(width=width, layers=layers, vocab=2000, context=77, embed_dim=640)
    model = _clip(cfg)
    toks = synth.make_tokens(n, 2000, 77, seed=5)
    ref = O.clip_encode_text(toks, synth.make_text_state_dict(**cfg), heads=width // 64)
    assert rel(model.encode_text(toks.cuda()), ref) < 1e-5


def test_classifier_head_init_from_text_encoder():
    """Classifier.init_head_with_text_encoder == FS/kadaptation_clip.py:136-138 (+ :154-155 logit scale)."""
    from peft_b200.classifier import Classifier
    cfg = dict(width=512, layers=1, vocab=1000, context=77, embed_dim=512)
    model = _clip(cfg)
    toks = synth.make_tokens(20, 1000, 77, seed=9)
    z = model.zeroshot_head_weights([toks[2 * i:2 * i + 2].cuda() for i in range(10)])
    clf = Classifier(model.visual, 512, 10).cuda()
    clf.init_head_with_text_encoder(z, logit_scale=model.logit_scale.detach())
    ref = O.zeroshot_weights([toks[2 * i:2 * i + 2] for i in range(10)], synth.make_text_state_dict(**cfg)) * float(model.logit_scale.exp())
    assert rel(clf.layers[0].weight, ref) < 1e-5
    assert float(clf.layers[0].bias.abs().max()) == 0.0
    with pytest.raises(RuntimeError):
        model.encode_text(toks)            # CPU tokens: no fallback
