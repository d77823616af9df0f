"""Flow-matching loss (CFM.forward, reference cfm.py:231-302) through f5b_cfm_loss vs the golden results of the
UNMODIFIED reference (tests/golden/loss_*.npz, oracle/make_golden.py) under the same replayed random draws."""
import numpy as np
import pytest
import torch

from helpers import O, build_cfm, max_abs
from test_oracle_golden import LOSS_CASES, loss_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", LOSS_CASES)
def test_loss_vs_reference_golden(name, golden_dir, precision="fp32", tol=1e-3, ltol=1e-4):
    cfg, sd, x1, text, lens, gold = loss_case(name, golden_dir)
    d = O.draw_loss_randoms(x1, int(gold["seed"]))
    cfm = build_cfm(cfg, sd, precision)
    loss, cond, pred = cfm(x1.cuda(), text.cuda(), lens=lens.cuda(),
                           draws=dict(frac_lengths=d.frac_lengths, rand=d.rand, x0=d.x0, time=d.time,
                                      drop_audio_cond=d.drop_audio_cond, drop_text=d.drop_text))
    assert max_abs(cond, gold["cond"]) == 0.0  # span mask + select: exact
    keep = O.lens_to_mask(lens, length=x1.shape[1])[..., None]  # frames past an utterance's length are not compared
    err = max_abs(pred.cpu() * keep, torch.as_tensor(gold["pred"]) * keep)
    rel = abs(float(loss) - float(gold["loss"])) / float(gold["loss"])
    print(f"{name} {precision}: pred max-abs {err:.3e}, loss {float(loss):.6f} vs {float(gold['loss']):.6f} (rel {rel:.2e})")
    assert err <= tol and rel <= ltol


@pytest.mark.parametrize("backbone,precision,tol,ltol", [("DiT", "fp32", 1e-3, 1e-4), ("DiT", "bf16", 2e-2, 5e-3),
                                                         ("UNetT", "bf16", 2e-2, 5e-3)])
def test_loss_vs_oracle_model_width(backbone, precision, tol, ltol):
    """dim 1024 (the width the 16-bit tensor-core mode needs), ragged batch, against the oracle restatement (which the
    CPU suite pins to the reference's golden loss vectors)."""
    if backbone == "DiT":
        cfg = O.ArchCfg.named("F5TTS_v1_Base", depth=2, attn_mask_enabled=True)
    else:
        cfg = O.ArchCfg.named("E2TTS_Base", depth=2)
    sd = O.synth_backbone_weights(cfg, seed=8)
    x1, text = O.synth_inputs(200, 200, 40, batch=3, seed=12)
    lens = torch.tensor([200, 133, 171])
    d = O.draw_loss_randoms(x1, 31)
    ref_loss, ref_cond, ref_pred = O.cfm_loss(sd, cfg, x1, text, lens, d)
    cfm = build_cfm(cfg, sd, precision)
    loss, cond, pred = cfm(x1.cuda(), text.cuda(), lens=lens.cuda(),
                           draws=dict(frac_lengths=d.frac_lengths, rand=d.rand, x0=d.x0, time=d.time,
                                      drop_audio_cond=d.drop_audio_cond, drop_text=d.drop_text))
    assert max_abs(cond, ref_cond) == 0.0
    keep = O.lens_to_mask(lens, length=x1.shape[1])[..., None]
    err = max_abs(pred.cpu() * keep, ref_pred * keep)
    rel = abs(float(loss) - float(ref_loss)) / float(ref_loss)
    print(f"{backbone} {precision}: pred max-abs {err:.3e}, loss {float(loss):.6f} vs {float(ref_loss):.6f} (rel {rel:.2e})")
    assert err <= tol and rel <= ltol


def test_loss_seeded_draws_are_the_reference_sequence():
    """Without injected draws the mirror draws in the reference's order: on a CPU-seeded generator moved to the device
    the span mask it builds equals the oracle's for the same numbers, the loss is finite and repeatable."""
    cfg = O.ArchCfg(backbone="DiT", dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2)
    sd = O.synth_backbone_weights(cfg, seed=3)
    cond, text = O.synth_inputs(60, 60, 12, batch=2, seed=4)
    lens = torch.tensor([60, 41])
    cfm = build_cfm(cfg, sd, "fp32")
    import random

    out = []
    for _ in range(2):
        torch.manual_seed(7)
        torch.cuda.manual_seed(7)
        random.seed(7)
        out.append(cfm(cond.cuda(), text.cuda(), lens=lens.cuda()))
    assert torch.isfinite(out[0][0]) and torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][2], out[1][2])
    # cond is x1 outside one contiguous span per utterance of 70-100 % of its length (frac_lengths_mask)
    zeroed = (out[0][1] == 0).all(dim=-1).cpu()
    for b, n in enumerate(lens.tolist()):
        z = zeroed[b, :n].nonzero().flatten()
        assert 0.7 * n - 1 <= len(z) <= n and (len(z) == 0 or int(z[-1] - z[0]) + 1 == len(z))
