"""Small end-to-end case for compute-sanitizer (tools/sanitize.sh): every kernel family of the path once, at sizes the
sanitizer finishes in a minute or two.  CUDA graphs are switched off by the caller (F5B_NO_GRAPH=1: the sanitizer sees
plain launches), programmatic dependent launch stays ON (that is what the racecheck is for)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "korean-f5-tts_b200")):
    sys.path.insert(0, p)

import torch  # noqa: E402

import f5b200  # noqa: E402
from oracle import f5_oracle as O  # noqa: E402  (inputs + reference values only)


def build(cfg, sd, precision):
    kw = dict(dim=cfg.dim, depth=cfg.depth, heads=cfg.heads, ff_mult=cfg.ff_mult, text_dim=cfg.text_dim,
              conv_layers=cfg.conv_layers, text_num_embeds=cfg.text_num_embeds, pe_attn_head=cfg.pe_attn_head,
              text_mask_padding=cfg.text_mask_padding, attn_mask_enabled=cfg.attn_mask_enabled, precision=precision)
    m = (f5b200.DiT if cfg.backbone == "DiT" else f5b200.UNetT)(**kw)
    m.load_state_dict(sd)
    return f5b200.CFM(transformer=m.cuda(), mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    torch.manual_seed(0)
    if which in ("all", "fp32"):
        cfg = O.ArchCfg(dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2, attn_mask_enabled=True)
        sd = O.synth_backbone_weights(cfg, seed=1)
        cond, text = O.synth_inputs(96, 30, 12, seed=2, batch=2)
        durs, lens = torch.tensor([96, 71]), torch.tensor([30, 22])
        ref = O.cfm_sample(sd, cfg, cond, text, durs, lens=lens, steps=3, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=3)
        out, _ = build(cfg, sd, "fp32").sample(cond=cond.cuda(), text=text.cuda(), duration=durs.cuda(), lens=lens.cuda(),
                                               steps=3, cfg_strength=2.0, sway_sampling_coef=-1.0, y0=ref.y0)
        keep = O.lens_to_mask(durs)[..., None]
        print("fp32 dim128 B=2 masked: max-abs", float(((out.cpu() - ref.out) * keep).abs().max()), flush=True)
    if which in ("all", "bf16"):
        cfg = O.ArchCfg.named("F5TTS_v1_Base", depth=2, attn_mask_enabled=True)
        sd = O.synth_backbone_weights(cfg, seed=4)
        cond, text = O.synth_inputs(200, 60, 30, seed=5, batch=2)
        durs, lens = torch.tensor([200, 150]), torch.tensor([60, 45])
        ref = O.cfm_sample(sd, cfg, cond, text, durs, lens=lens, steps=2, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=6)
        out, _ = build(cfg, sd, "bf16").sample(cond=cond.cuda(), text=text.cuda(), duration=durs.cuda(), lens=lens.cuda(),
                                               steps=2, cfg_strength=2.0, sway_sampling_coef=-1.0, y0=ref.y0)
        keep = O.lens_to_mask(durs)[..., None]
        print("bf16 dim1024 depth2 B=2 masked: max-abs", float(((out.cpu() - ref.out) * keep).abs().max()), flush=True)
    if which in ("all", "unett"):
        cfg = O.ArchCfg.named("E2TTS_Base", depth=2)
        sd = O.synth_backbone_weights(cfg, seed=7)
        cond, text = O.synth_inputs(150, 40, 30, seed=8)
        ref = O.cfm_sample(sd, cfg, cond, text, 150, steps=2, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=9)
        out, _ = build(cfg, sd, "bf16").sample(cond=cond.cuda(), text=text.cuda(), duration=150, steps=2, cfg_strength=2.0,
                                               sway_sampling_coef=-1.0, y0=ref.y0)
        print("bf16 UNetT depth2: max-abs", float((out.cpu() - ref.out).abs().max()), flush=True)
    if which in ("all", "vocos"):
        vsd = O.synth_vocos_weights(0)
        mel = (torch.randn(2, 100, 40) * 2 - 3)
        want = O.vocos_decode(vsd, mel)
        for precision in ("fp32", "bf16"):
            voc = f5b200.Vocos(precision=precision)
            voc.load_state_dict(vsd)
            wav = voc.cuda().decode(mel.cuda())
            print(f"vocos {precision}: max-abs", float((wav.cpu() - want).abs().max()), flush=True)
        wav = torch.randn(1, 24000) * 0.1
        m = f5b200.MelSpec(mel_spec_type="vocos").cuda()(wav.cuda())
        print("mel: max-abs", float((m.cpu() - O.mel_spectrogram(wav)).abs().max()), flush=True)
    print("sanitize_case done", flush=True)


if __name__ == "__main__":
    main()
