"""Checkpoint ingestion with the reference's signatures (src/f5_tts/infer/utils_infer.py:198-345): `.pt` /
`.safetensors` files as the reference's trainer writes them (EMA prefix, `initted` / `step` counters, legacy mel-spec
buffers) and PEFT / LoRA adapters merged into plain linears, loaded strictly into the f5b200 module tree (same parameter
names and shapes as the reference's).  Host-side tensor bookkeeping only; the packed 16-bit tensor-core copies of the
weights are made by the library when the plan is first used (`f5b_plan_create`)."""
from __future__ import annotations

import torch

from .backbones import DiT, UNetT  # noqa: F401  (model_cls candidates)
from .cfm import CFM

# utils_infer.py:58-69
target_sample_rate = 24000
n_mel_channels = 100
hop_length = 256
win_length = 1024
n_fft = 1024
mel_spec_type = "vocos"
ode_method = "euler"


def get_tokenizer(dataset_name: str, tokenizer: str = "custom"):
    """The "custom" branch of model/utils.py:142-149 (a vocab.txt path: one symbol per line, id = line number) and the
    "byte" branch (:138-140).  The dataset-relative branches need the reference's data directory and stay there."""
    if tokenizer == "byte":
        return None, 256
    if tokenizer != "custom":
        raise ValueError("pass the vocab.txt path with tokenizer='custom' (what load_model does, utils_infer.py:313)")
    with open(dataset_name, "r", encoding="utf-8") as f:
        vocab_char_map = {}
        for i, char in enumerate(f):
            vocab_char_map[char[:-1]] = i
    return vocab_char_map, len(vocab_char_map)


def convert_peft_state_dict_to_plain(state_dict: dict, lora_alpha: float = 32.0, lora_r: int = 16) -> dict:
    """utils_infer.py:198-239: strip `base_model.model.`, fold `lora_B @ lora_A * alpha / r` into `base_layer.weight`,
    rename `base_layer.{weight,bias}` to the plain names.  A plain state dict passes through untouched."""
    prefix = "base_model.model."
    if not any(k.startswith(prefix) for k in state_dict):
        return state_dict
    scale = lora_alpha / lora_r
    plain = {k[len(prefix):]: v for k, v in state_dict.items() if k.startswith(prefix)}
    drop = set()
    for k in list(plain):
        if k.endswith(".base_layer.weight"):
            stem = k[: -len(".base_layer.weight")]
            a_key, b_key = stem + ".lora_A.default.weight", stem + ".lora_B.default.weight"
            if a_key in plain and b_key in plain:
                w = plain[k]
                plain[stem + ".weight"] = (w + (plain[b_key] @ plain[a_key]).to(w.dtype) * scale).to(w.dtype)
                drop.update((k, a_key, b_key))
        elif k.endswith(".base_layer.bias"):
            plain.setdefault(k.replace(".base_layer.bias", ".bias"), plain[k])
            drop.add(k)
    for k in drop:
        plain.pop(k, None)
    return plain


def load_checkpoint(model, ckpt_path: str, device: str, dtype=None, use_ema=True):
    """utils_infer.py:242-286.  dtype None on a CUDA device selects the 16-bit tensor-core mode (the reference picks
    fp16 there); torch.float32 keeps the fp32 kernels.  Parameters stay fp32 masters either way."""
    if dtype is None:
        dtype = torch.float16 if "cuda" in str(device) else torch.float32
    model.transformer.to(dtype)  # selects the kernel set; the parameters stay fp32 masters (backbones.py `to`)
    ckpt_type = ckpt_path.split(".")[-1]
    if ckpt_type == "safetensors":
        from safetensors.torch import load_file

        checkpoint = load_file(ckpt_path, device="cpu")
    else:
        checkpoint = torch.load(ckpt_path, map_location="cpu", weights_only=True)
    if use_ema:
        if ckpt_type == "safetensors":
            checkpoint = {"ema_model_state_dict": checkpoint}
        state = {k.replace("ema_model.", ""): v for k, v in checkpoint["ema_model_state_dict"].items()
                 if k not in ["initted", "step"]}
        for key in ["mel_spec.mel_stft.mel_scale.fb", "mel_spec.mel_stft.spectrogram.window"]:
            state.pop(key, None)
    else:
        if ckpt_type == "safetensors":
            checkpoint = {"model_state_dict": checkpoint}
        state = checkpoint["model_state_dict"]
    model.load_state_dict(convert_peft_state_dict_to_plain(state))
    return model.to(device)


def load_model(model_cls, model_cfg, ckpt_path, mel_spec_type=mel_spec_type, vocab_file="", ode_method=ode_method,
               use_ema=True, device="cuda", use_skip_tc=False, use_n2gk_plus=True, tokenizer="custom"):
    """utils_infer.py:292-345: vocab -> `text_num_embeds = vocab_size + 1`, CFM(model_cls(**model_cfg)), checkpoint,
    and the three tokenizer attributes `infer_batch_process` reads back (:550-553)."""
    if vocab_file == "":
        raise ValueError("vocab_file is required (the reference's bundled infer/examples/vocab.txt is not shipped here)")
    if mel_spec_type != "vocos":
        raise ValueError("only the Vocos mel / vocoder pair is built (BigVGAN is out of scope)")
    vocab_char_map, vocab_size = get_tokenizer(vocab_file, "custom")
    model = CFM(
        transformer=model_cls(**model_cfg, text_num_embeds=vocab_size + 1, mel_dim=n_mel_channels),
        mel_spec_kwargs=dict(n_fft=n_fft, hop_length=hop_length, win_length=win_length, n_mel_channels=n_mel_channels,
                             target_sample_rate=target_sample_rate, mel_spec_type=mel_spec_type),
        odeint_kwargs=dict(method=ode_method),
        vocab_char_map=vocab_char_map,
    )
    model = load_checkpoint(model, ckpt_path, device, dtype=None, use_ema=use_ema)
    model._use_skip_tc = use_skip_tc
    model._use_n2gk_plus = use_n2gk_plus
    model._tokenizer_type = tokenizer
    return model
