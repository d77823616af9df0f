"""Checkpoint ingestion (SURVEY.md 8b `load_model` / 8f row f3): files laid out as the reference's trainer and PEFT
write them load strictly into the f5b200 tree.  CPU: formats, EMA prefix, LoRA merge against the restated formula
(infer/utils_infer.py:198-286).  GPU: a model loaded from a LoRA `.safetensors` samples like the merged weights."""
import os

import pytest
import torch

import f5b200
from f5b200 import checkpoint as CK
from helpers import O, max_abs

CFG = dict(dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2)
VOCAB = [" ", "a", "b", "c", "d", ",", "."]


def _vocab_file(tmp_path):
    p = tmp_path / "vocab.txt"
    p.write_text("".join(c + "\n" for c in VOCAB), encoding="utf-8")
    return str(p)


def _plain_state(seed=3):
    cfg = O.ArchCfg(**CFG, text_num_embeds=len(VOCAB) + 1)
    return cfg, {"transformer." + k: v for k, v in O.synth_backbone_weights(cfg, seed=seed).items()}


def _as_lora(state, r=4, seed=0):
    """PEFT layout of the same model: attention linears split into base_layer + lora_A / lora_B (utils_infer.py:198)."""
    g = torch.Generator().manual_seed(seed)
    out, merged = {}, dict(state)
    for k, v in state.items():
        stem = k[: -len(".weight")] if k.endswith(".weight") else None
        if stem and stem.endswith((".to_q", ".to_k", ".to_v", ".to_out.0")):
            a = torch.randn(r, v.shape[1], generator=g) * 0.05
            b = torch.randn(v.shape[0], r, generator=g) * 0.05
            out["base_model.model." + stem + ".base_layer.weight"] = v
            out["base_model.model." + stem + ".lora_A.default.weight"] = a
            out["base_model.model." + stem + ".lora_B.default.weight"] = b
            merged[k] = v + (b @ a) * (32.0 / 16)
        elif k.endswith(".bias") and k[: -len(".bias")].endswith((".to_q", ".to_k", ".to_v", ".to_out.0")):
            out["base_model.model." + k[: -len(".bias")] + ".base_layer.bias"] = v
        else:
            out["base_model.model." + k] = v
    return out, merged


def test_get_tokenizer_custom(tmp_path):
    m, n = CK.get_tokenizer(_vocab_file(tmp_path), "custom")
    assert n == len(VOCAB) and m[" "] == 0 and m["."] == 6
    assert CK.get_tokenizer("", "byte") == (None, 256)


def test_lora_merge_formula():
    _, state = _plain_state()
    lora, merged = _as_lora(state)
    plain = CK.convert_peft_state_dict_to_plain(lora)
    assert sorted(plain) == sorted(merged)
    for k in merged:
        assert torch.equal(plain[k], merged[k]), k
    assert CK.convert_peft_state_dict_to_plain(state) is state  # plain dicts pass through


@pytest.mark.parametrize("fmt,use_ema", [("pt", True), ("pt", False), ("safetensors", True), ("safetensors", False)])
def test_load_checkpoint_formats_cpu(tmp_path, fmt, use_ema):
    """No GPU needed: loading is host-side bookkeeping (device='cpu' keeps the module on the host; computing there
    still raises)."""
    cfg, state = _plain_state()
    path = str(tmp_path / f"model.{fmt}")
    if fmt == "pt":
        if use_ema:
            ema = {"ema_model." + k: v for k, v in state.items()}
            ema.update(initted=torch.tensor(True), step=torch.tensor(7))
            ema["ema_model.mel_spec.mel_stft.mel_scale.fb"] = torch.zeros(3)  # legacy buffer, dropped (:270-273)
            torch.save({"ema_model_state_dict": ema}, path)
        else:
            torch.save({"model_state_dict": state}, path)
    else:
        from safetensors.torch import save_file

        save_file({("ema_model." + k if use_ema else k): v.contiguous() for k, v in state.items()}, path)
    model = f5b200.CFM(transformer=f5b200.DiT(**CFG, text_num_embeds=len(VOCAB) + 1, mel_dim=100),
                       mel_spec_kwargs=dict(mel_spec_type="vocos"))
    model = CK.load_checkpoint(model, path, "cpu", use_ema=use_ema)
    got = model.state_dict()
    for k, v in state.items():
        assert torch.equal(got[k], v), k
    assert model.transformer.precision == "fp32"


@pytest.mark.parametrize("with_inv_freq", [True, False])
def test_reference_checkpoints_carry_rotary_inv_freq(tmp_path, with_inv_freq):
    """trainer-written checkpoints hold `transformer.rotary_embed.inv_freq` (x_transformers registers
    it as a persistent buffer); a strict load must accept it -- and still accept a file written without it."""
    cfg, state = _plain_state()
    ema = {"ema_model." + k: v for k, v in state.items()}
    if with_inv_freq:
        ema["ema_model.transformer.rotary_embed.inv_freq"] = 1.0 / (10000.0 ** (torch.arange(0, 64, 2).float() / 64))
    ema.update(initted=torch.tensor(True), step=torch.tensor(7))
    path = str(tmp_path / "model.pt")
    torch.save({"ema_model_state_dict": ema}, path)
    model = f5b200.CFM(transformer=f5b200.DiT(**CFG, text_num_embeds=len(VOCAB) + 1, mel_dim=100),
                       mel_spec_kwargs=dict(mel_spec_type="vocos"))
    model = CK.load_checkpoint(model, path, "cpu", use_ema=True)
    assert "transformer.rotary_embed.inv_freq" in model.state_dict()


def test_live_reference_state_dict_loads_strictly(tmp_path):
    """The state dict of the reference's own CFM(DiT) (unmodified modules through the loader) loads strictly."""
    from oracle import make_golden, ref_loader

    if not ref_loader.available():
        pytest.skip("/root/reference is only present in the build container")
    R = ref_loader.load()
    cfg, state = _plain_state()
    ref = make_golden.build_reference(R, cfg, {k[len("transformer."):]: v for k, v in state.items()})
    sd = {k: v for k, v in ref.state_dict().items() if not k.startswith("mel_spec.")}
    assert "transformer.rotary_embed.inv_freq" in sd
    torch.save({"model_state_dict": sd}, str(tmp_path / "ref.pt"))
    model = f5b200.CFM(transformer=f5b200.DiT(**CFG, text_num_embeds=len(VOCAB) + 1, mel_dim=100),
                       mel_spec_kwargs=dict(mel_spec_type="vocos"))
    CK.load_checkpoint(model, str(tmp_path / "ref.pt"), "cpu", use_ema=False)
    for k, v in sd.items():
        assert torch.equal(model.state_dict()[k], v), k


def test_load_checkpoint_is_strict(tmp_path):
    _, state = _plain_state()
    state.pop(next(iter(state)))
    path = str(tmp_path / "model.pt")
    torch.save({"model_state_dict": state}, path)
    model = f5b200.CFM(transformer=f5b200.DiT(**CFG, text_num_embeds=len(VOCAB) + 1, mel_dim=100),
                       mel_spec_kwargs=dict(mel_spec_type="vocos"))
    with pytest.raises(RuntimeError):
        CK.load_checkpoint(model, path, "cpu", use_ema=False)


@pytest.mark.gpu
def test_load_model_lora_safetensors_samples_like_merged(tmp_path):
    from safetensors.torch import save_file

    cfg, state = _plain_state()
    lora, merged = _as_lora(state)
    path = str(tmp_path / "model_last.safetensors")
    save_file({"ema_model." + k: v.contiguous() for k, v in lora.items()}, path)
    model = f5b200.load_model(f5b200.DiT, CFG, path, vocab_file=_vocab_file(tmp_path), device="cuda", tokenizer="kor_phoneme")
    assert (model._use_skip_tc, model._use_n2gk_plus, model._tokenizer_type) == (False, True, "kor_phoneme")
    assert model.transformer.precision == "bf16" and next(model.parameters()).dtype == torch.float32
    model.transformer.float()  # compare in the fp32 mode
    cond, _ = O.synth_inputs(90, 30, 8, seed=4)
    text = [list("abc, dab."), ]
    y0 = torch.randn(1, 90, 100, generator=torch.Generator().manual_seed(1))
    out, _ = model.sample(cond=cond.cuda(), text=text, duration=90, steps=4, cfg_strength=2.0, sway_sampling_coef=-1.0, y0=y0)
    sd = {k[len("transformer."):]: v for k, v in merged.items()}
    ref = O.cfm_sample(sd, cfg, cond, text, 90, steps=4, cfg_strength=2.0, sway_sampling_coef=-1.0, y0=y0,
                       vocab={c: i for i, c in enumerate(VOCAB)})
    err = max_abs(out.cpu(), ref.out)
    print(f"LoRA checkpoint -> sample vs oracle on merged weights: {err:.2e}")
    assert err <= 1e-3
