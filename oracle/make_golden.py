"""Generate tests/golden/*.npz by running the UNMODIFIED reference modules (container only).

TEST INFRASTRUCTURE.  Run:  python -m oracle.make_golden
The reference holds no golden vectors for this path, so these fixtures are produced by
importing the reference's own cfm.py / dit.py / unett.py / modules.py (oracle/ref_loader.py) on seeded
synthetic weights (oracle.f5_oracle.synth_backbone_weights) and seeded inputs.  Weights are re-synthesised
from the seed at test time; a checksum is stored to detect drift.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import f5_oracle as O
from . import ref_loader

CASES = {
    # name: (ArchCfg kwargs, batch, lens, durations, n_text, steps, cfg_strength, sway)
    "dit_v1style_b1": (dict(backbone="DiT", dim=128, depth=3, heads=2, ff_mult=2, text_dim=64, conv_layers=2,
                            pe_attn_head=None, text_mask_padding=True), 1, [30], [90], 14, 16, 2.0, -1.0),
    "dit_v1style_b3_mask": (dict(backbone="DiT", dim=128, depth=3, heads=2, ff_mult=2, text_dim=64, conv_layers=2,
                                 pe_attn_head=None, text_mask_padding=True, attn_mask_enabled=True),
                            3, [30, 22, 17], [90, 61, 75], 14, 7, 2.0, -1.0),
    "dit_smallstyle_b2_nomask": (dict(backbone="DiT", dim=128, depth=2, heads=2, ff_mult=2, text_dim=64,
                                      conv_layers=2, pe_attn_head=1, text_mask_padding=False),
                                 2, [28, 20], [70, 64], 11, 32, 2.0, -1.0),
    "unett_b1": (dict(backbone="UNetT", dim=128, depth=4, heads=2, ff_mult=4, text_dim=100, conv_layers=0,
                      pe_attn_head=1, text_mask_padding=False), 1, [25], [80], 12, 10, 2.0, None),
    "dit_nocfg_b1": (dict(backbone="DiT", dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2,
                          pe_attn_head=None, text_mask_padding=True), 1, [30], [77], 9, 5, 0.0, -1.0),
}


def build_reference(R, cfg: O.ArchCfg, sd):
    kw = dict(dim=cfg.dim, depth=cfg.depth, heads=cfg.heads, ff_mult=cfg.ff_mult, text_dim=cfg.text_dim,
              conv_layers=cfg.conv_layers, text_num_embeds=cfg.text_num_embeds, pe_attn_head=cfg.pe_attn_head,
              text_mask_padding=cfg.text_mask_padding, attn_mask_enabled=cfg.attn_mask_enabled, qk_norm=cfg.qk_norm)
    m = (R.DiT if cfg.backbone == "DiT" else R.UNetT)(**kw)
    # x_transformers registers RotaryEmbedding.inv_freq as a persistent buffer (so does the shim): synthetic state dicts
    # do not carry it, real checkpoints do
    m.load_state_dict({**sd, "rotary_embed.inv_freq": m.rotary_embed.inv_freq}, strict=True)
    return R.CFM(transformer=m, mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()


def case_inputs(name):
    kw, B, lens, durs, nt, steps, cfgs, sway = CASES[name]
    cfg = O.ArchCfg(**kw)
    cond, text = O.synth_inputs(max(durs), max(lens), nt, batch=B, seed=len(name))
    if B > 1:
        text[1, nt - 4:] = -1
    return cfg, cond, text, torch.tensor(lens), torch.tensor(durs), steps, cfgs, sway


def main():
    R = ref_loader.load()
    out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name in CASES:
        cfg, cond, text, lens, durs, steps, cfgs, sway = case_inputs(name)
        sd = O.synth_backbone_weights(cfg, seed=11)
        cfm = build_reference(R, cfg, sd)
        out, traj = cfm.sample(cond=cond, text=text, duration=durs, lens=lens, steps=steps, cfg_strength=cfgs,
                               sway_sampling_coef=sway, seed=5)
        wsum = float(sum(v.double().abs().sum() for v in sd.values()))
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), out=out.numpy(), traj_first=traj[1].numpy(),
                            traj_mid=traj[steps // 2].numpy(), weight_abs_sum=np.float64(wsum))
        print(name, tuple(out.shape), "weights |sum|", wsum)
    # flow-matching loss (CFM.forward, cfm.py:231-302): the reference run under seeded CPU generators
    import random
    for name, seed in (("dit_v1style_b3_mask", 21), ("unett_b1", 22), ("dit_smallstyle_b2_nomask", 23)):
        cfg, cond, text, lens, durs, *_ = case_inputs(name)
        sd = O.synth_backbone_weights(cfg, seed=11)
        cfm = build_reference(R, cfg, sd)
        x1 = cond[:, : int(lens.max())]  # the "training mel" batch: the prompt frames, ragged by lens
        torch.manual_seed(seed)
        random.seed(seed)
        with torch.no_grad():
            loss, c, pred = cfm(x1, text, lens=lens)
        np.savez_compressed(os.path.join(out_dir, "loss_" + name + ".npz"), loss=loss.numpy(), cond=c.numpy(),
                            pred=pred.numpy(), seed=np.int64(seed))
        print("loss", name, float(loss))
    # mel frontend: reference MelSpec on a seeded waveform
    g = torch.Generator().manual_seed(3)
    wav = torch.randn(2, 24000, generator=g) * 0.1
    mel = R.modules.MelSpec()(wav)
    np.savez_compressed(os.path.join(out_dir, "mel_frontend.npz"), mel=mel.numpy())
    # host-side integer helpers
    t16 = R.utils.get_epss_timesteps(16, device="cpu", dtype=torch.float32)
    t16 = t16 + (-1.0) * (torch.cos(torch.pi / 2 * t16) - 1 + t16)
    np.savez_compressed(os.path.join(out_dir, "host_helpers.npz"), t16_sway=t16.numpy(),
                        mask=R.utils.lens_to_mask(torch.tensor([3, 5, 1])).numpy(),
                        idx=R.utils.list_str_to_idx([["a", "b", "?"], ["b"]], {"a": 1, "b": 2}).numpy())
    print("done")


if __name__ == "__main__":
    main()
