"""The backward pass (reference trainer.py:384 `accelerator.backward(loss)` over CFM.forward,
cfm.py:231-302): gradients of EVERY parameter of the DiT (blocks, norm_out, proj_out, timestep MLP, input embedding with
its grouped convolutions, text embedding with its ConvNeXt-V2 / GRN blocks) and of the residual stream entering the last
block / block 0, through `f5b_cfm_loss_tail_grads`.
  * against golden gradients of the UNMODIFIED reference (tests/golden/grads_*.npz, oracle/make_golden_grads.py);
  * at model width 1024 against torch.autograd over the oracle, in the fp32 and the 16-bit forward modes."""
import os

import numpy as np
import pytest
import torch

from helpers import O, build_cfm, max_abs
from test_oracle_golden import loss_case

pytestmark = pytest.mark.gpu
GRAD_CASES = ["dit_v1style_b3_mask", "dit_smallstyle_b2_nomask"]


def _draws(d):
    return dict(frac_lengths=d.frac_lengths, rand=d.rand, x0=d.x0, time=d.time, drop_audio_cond=d.drop_audio_cond,
                drop_text=d.drop_text)


@pytest.mark.parametrize("name", GRAD_CASES)
def test_tail_gradients_vs_reference_golden(name, golden_dir):
    cfg, sd, x1, text, lens, _ = loss_case(name, golden_dir)
    gold = np.load(os.path.join(golden_dir, f"grads_{name}.npz"))
    d = O.draw_loss_randoms(x1, int(gold["seed"]))
    cfm = build_cfm(cfg, sd, "fp32")
    loss, grads = cfm.tail_gradients(x1.cuda(), text.cuda(), lens=lens.cuda(), draws=_draws(d))
    assert abs(float(loss) - float(gold["loss"])) <= 1e-4 * float(gold["loss"])
    names = [k for k in gold.files if k not in ("loss", "seed")]
    # last block + norm_out + proj_out + d_block_input in full; earlier blocks, time_embed, d_input_embed every 5th element
    assert {k.split(":")[-1] for k in names} == set(grads)  # every DiT parameter + the two stream gradients
    assert len(names) == 14 * cfg.depth + 4 + 4 + 6 + 1 + 10 * cfg.conv_layers + 2
    keep = O.lens_to_mask(lens, length=x1.shape[1])[..., None]
    worst = 0.0
    for k in names:
        sub = int(k.split(":")[0][3:]) if k.startswith("sub") else 1
        g, w = grads[k.split(":")[-1]].cpu(), torch.from_numpy(gold[k])
        if k.endswith(("d_block_input", "d_input_embed")):  # frames past an utterance's length carry no defined gradient
            g = g * keep
            w = w * keep if sub == 1 else w * keep.expand_as(g).flatten()[::sub]
        g = g.flatten()[::sub].reshape(w.shape)
        rel = max_abs(g, w) / max(float(w.abs().max()), 1e-8)
        worst = max(worst, rel)
        assert rel <= 1e-3, (k, rel)
    print(f"{name}: {len(names)} gradients vs the reference's autograd, worst max-abs / max|g| = {worst:.2e}")


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-3), ("bf16", 3e-2)])
def test_tail_gradients_vs_oracle_autograd_width_1024(precision, tol):
    """dim 1024, ragged masked batch: torch.autograd over the oracle's cfm_loss (weights as leaves).  In the 16-bit mode
    only the forward evaluation feeding the block is 16-bit; the tail is differentiated in fp32."""
    cfg = O.ArchCfg.named("F5TTS_v1_Base", depth=2, attn_mask_enabled=True)
    sd = O.synth_backbone_weights(cfg, seed=8)
    x1, text = O.synth_inputs(160, 160, 30, batch=2, seed=12)
    lens = torch.tensor([160, 117])
    d = O.draw_loss_randoms(x1, 31)
    leaves = {k: v.clone().requires_grad_(True) for k, v in sd.items()}  # every parameter of the DiT
    ref_loss, _, _ = O.cfm_loss({**sd, **leaves}, cfg, x1, text, lens, d)
    ref_loss.backward()
    cfm = build_cfm(cfg, sd, precision)
    loss, grads = cfm.tail_gradients(x1.cuda(), text.cuda(), lens=lens.cuda(), draws=_draws(d))
    assert abs(float(loss) - float(ref_loss)) <= (1e-4 if precision == "fp32" else 5e-3) * float(ref_loss)
    worst = 0.0
    for k, leaf in leaves.items():
        g = grads["transformer." + k].cpu()
        rel = max_abs(g, leaf.grad) / max(float(leaf.grad.abs().max()), 1e-8)
        worst = max(worst, rel)
        assert rel <= tol, (k, rel)
    print(f"width 1024 {precision}: {len(leaves)} gradients vs autograd over the oracle, worst relative max-abs {worst:.2e}")


def test_unknown_gradient_name_is_loud():
    cfg = O.ArchCfg(dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2)
    sd = O.synth_backbone_weights(cfg, seed=3)
    cfm = build_cfm(cfg, sd, "fp32")
    x1, text = O.synth_inputs(40, 40, 8, seed=1)
    with pytest.raises(KeyError):  # a name that is no parameter must not come back as zeros
        cfm.tail_gradients(x1.cuda(), text.cuda(), names=["input_embed.proj.weights"])


def test_train_step_matches_torch_adamw_on_the_oracle():
    """Two optimiser steps of `TrainStep` (backward + global-norm clip + AdamW, flat buffers) against torch.optim.AdamW +
    clip_grad_norm_ driven by autograd over the oracle with the same draws (reference trainer.py:146,384-390)."""
    import f5b200

    cfg = O.ArchCfg(dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2, attn_mask_enabled=True)
    sd = O.synth_backbone_weights(cfg, seed=5)
    x1, text = O.synth_inputs(60, 60, 12, batch=2, seed=3)
    lens = torch.tensor([60, 44])
    cfm = build_cfm(cfg, sd, "fp32")
    ts = f5b200.TrainStep(cfm, lr=1e-3, weight_decay=1e-2, max_grad_norm=0.05, ema_decay=0.9)
    leaves = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    opt = torch.optim.AdamW(list(leaves.values()), lr=1e-3, weight_decay=1e-2)
    ema = {k: v.detach().clone() for k, v in leaves.items()}
    for it in range(2):
        d = O.draw_loss_randoms(x1, 40 + it)
        loss = ts.step(x1.cuda(), text.cuda(), lens=lens.cuda(), draws=_draws(d))
        opt.zero_grad()
        ref_loss, _, _ = O.cfm_loss(leaves, cfg, x1, text, lens, d)
        ref_loss.backward()
        norm = torch.nn.utils.clip_grad_norm_(list(leaves.values()), 0.05)
        opt.step()
        for k in ema:
            ema[k] = 0.9 * ema[k] + 0.1 * leaves[k].detach()
        assert abs(float(loss) - float(ref_loss.detach())) <= 1e-4 * float(ref_loss.detach())
        assert abs(ts.last["grad_norm"] - float(norm)) <= 1e-4 * float(norm) and ts.last["clip_coef"] < 1.0
    got = cfm.transformer.state_dict()
    worst = max(max_abs(got[k], v.detach()) / max(float(v.detach().abs().max()), 1e-8) for k, v in leaves.items())
    e = ts.ema_state_dict()
    worst_ema = max(max_abs(e["ema_model.transformer." + k], v) / max(float(v.abs().max()), 1e-8) for k, v in ema.items())
    print(f"two AdamW steps: parameters within {worst:.2e}, EMA within {worst_ema:.2e} (relative to max|w|) of torch.optim.AdamW")
    assert worst <= 1e-4 and worst_ema <= 1e-4


def test_step_workspace_survives_changing_batch_shapes():
    """The training step's workspace is kept by the plan and handed out again on the next call (grow-only): batches of
    changing shape -- what the reference's frame-budget batch sampler produces -- give the bits they give on a fresh
    model, whether the call before was larger or smaller."""
    cfg = O.ArchCfg(dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2, attn_mask_enabled=True)
    sd = O.synth_backbone_weights(cfg, seed=3)
    shapes = [(2, 48), (3, 80), (1, 33), (3, 80), (2, 48)]
    cases = {}
    for b, n in set(shapes):
        x1, text = O.synth_inputs(n, n, 9, batch=b, seed=b * 100 + n)
        lens = torch.tensor([n - 3 * i for i in range(b)])
        cases[(b, n)] = (x1.cuda(), text.cuda(), lens.cuda(), _draws(O.draw_loss_randoms(x1, 7)))
    fresh = {}
    for key, (x1, text, lens, d) in cases.items():
        loss, g = build_cfm(cfg, sd, "fp32").tail_gradients(x1, text, lens=lens, draws=d)
        fresh[key] = (loss.cpu(), {k: v.cpu() for k, v in g.items()})
    cfm = build_cfm(cfg, sd, "fp32")
    for key in shapes:
        x1, text, lens, d = cases[key]
        loss, g = cfm.tail_gradients(x1, text, lens=lens, draws=d)
        assert torch.equal(loss.cpu(), fresh[key][0])
        for k, v in g.items():
            assert torch.equal(v.cpu(), fresh[key][1][k]), (key, k)
