"""Shared test helpers: build f5b200 modules from the oracle's synthetic weights; error metrics."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "korean-f5-tts_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from oracle import f5_oracle as O  # noqa: E402  (checker only)


def build_backbone(cfg: "O.ArchCfg", sd: dict, precision: str, device="cuda"):
    import f5b200

    kw = dict(dim=cfg.dim, depth=cfg.depth, heads=cfg.heads, ff_mult=cfg.ff_mult, text_dim=cfg.text_dim,
              conv_layers=cfg.conv_layers, text_num_embeds=cfg.text_num_embeds, pe_attn_head=cfg.pe_attn_head,
              text_mask_padding=cfg.text_mask_padding, attn_mask_enabled=cfg.attn_mask_enabled, precision=precision)
    m = (f5b200.DiT if cfg.backbone == "DiT" else f5b200.UNetT)(**kw)
    m.load_state_dict(sd, strict=True)
    return m.to(device)


def build_cfm(cfg, sd, precision, device="cuda"):
    import f5b200

    return f5b200.CFM(transformer=build_backbone(cfg, sd, precision, device),
                      mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()


def max_abs(a, b) -> float:
    a = a.detach().cpu().double() if torch.is_tensor(a) else torch.as_tensor(np.asarray(a)).double()
    b = b.detach().cpu().double() if torch.is_tensor(b) else torch.as_tensor(np.asarray(b)).double()
    return float((a - b).abs().max())


def snr_db(ref, test) -> float:
    ref = ref.detach().cpu().double()
    test = test.detach().cpu().double()
    return float(10 * torch.log10(ref.pow(2).sum() / (ref - test).pow(2).sum().clamp_min(1e-300)))
