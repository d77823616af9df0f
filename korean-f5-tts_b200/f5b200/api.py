"""`F5TTS` -- the reference's user-facing class (src/f5_tts/api.py:23-149) over the B200 path: same constructor and
`infer` arguments and results `(wav, sr, spec)`.  Differences, all forced by the box (no network, no ASR / pydub):
`ckpt_file` and `vocab_file` are required (no hub download), the Vocos weights come from `vocoder_local_path` (or are
loaded afterwards into `.vocoder`), `transcribe` and `remove_silence` raise, `ref_file` may be a path to a PCM `.wav`
or an `(audio [channels, n], sample_rate)` pair.  All text chunks of a call run as variable-length batches
(`infer.infer_batch_process`)."""
from __future__ import annotations

import random
import re
import sys
import wave

import numpy as np
import torch

from . import backbones
from .checkpoint import load_model
from .infer import infer_batch_process
from .utils import seed_everything
from .vocoder import load_vocoder

# model.backbone / model.arch of the reference's configs/<name>.yaml
MODEL_CONFIGS = {
    "F5TTS_v1_Base": ("DiT", dict(dim=1024, depth=22, heads=16, ff_mult=2, text_dim=512, text_mask_padding=True,
                                  qk_norm=None, conv_layers=4, pe_attn_head=None, attn_backend="torch",
                                  attn_mask_enabled=False, checkpoint_activations=False)),
    "F5TTS_Base": ("DiT", dict(dim=1024, depth=22, heads=16, ff_mult=2, text_dim=512, text_mask_padding=False,
                               conv_layers=4, pe_attn_head=1, attn_backend="torch", attn_mask_enabled=False,
                               checkpoint_activations=False)),
    "F5TTS_Small": ("DiT", dict(dim=768, depth=18, heads=12, ff_mult=2, text_dim=512, text_mask_padding=False,
                                conv_layers=4, pe_attn_head=1, attn_backend="torch", attn_mask_enabled=False,
                                checkpoint_activations=False)),
    "E2TTS_Base": ("UNetT", dict(dim=1024, depth=24, heads=16, ff_mult=4, text_mask_padding=False, pe_attn_head=1)),
    "E2TTS_Small": ("UNetT", dict(dim=768, depth=20, heads=12, ff_mult=4, text_mask_padding=False, pe_attn_head=1)),
}


def model_config(model: str):
    """A name from the table above, or the path of one of the reference's YAML configs (model.backbone, model.arch)."""
    if model in MODEL_CONFIGS:
        backbone, arch = MODEL_CONFIGS[model]
        return backbone, dict(arch), "vocos", 24000
    import yaml

    with open(model, "r", encoding="utf-8") as f:
        m = yaml.safe_load(f)["model"]
    return m["backbone"], dict(m["arch"]), m["mel_spec"]["mel_spec_type"], m["mel_spec"]["target_sample_rate"]


def chunk_text(text: str, max_chars: int = 135):
    """Sentence-wise packing of the text into chunks of at most `max_chars` utf-8 bytes (utils_infer.py:83-110)."""
    chunks, current = [], ""
    for sentence in re.split(r"(?<=[;:,.!?])\s+|(?<=[；：，。！？])", text):
        piece = sentence + " " if sentence and len(sentence[-1].encode("utf-8")) == 1 else sentence
        if len(current.encode("utf-8")) + len(sentence.encode("utf-8")) <= max_chars:
            current += piece
        else:
            if current:
                chunks.append(current.strip())
            current = piece
    if current:
        chunks.append(current.strip())
    return chunks


def read_wav(path: str):
    """PCM `.wav` -> (float32 [channels, n] in [-1, 1), sample_rate); stands in for `torchaudio.load`
    (utils_infer.py:472), whose codecs are not on this image."""
    with wave.open(path, "rb") as w:
        ch, width, sr, n = w.getnchannels(), w.getsampwidth(), w.getframerate(), w.getnframes()
        raw = w.readframes(n)
    if width == 2:
        a = np.frombuffer(raw, dtype="<i2").astype(np.float32) / 32768.0
    elif width == 4:
        a = np.frombuffer(raw, dtype="<i4").astype(np.float32) / 2147483648.0
    elif width == 1:
        a = (np.frombuffer(raw, dtype=np.uint8).astype(np.float32) - 128.0) / 128.0
    else:
        raise ValueError(f"unsupported PCM sample width {width}")
    return torch.from_numpy(a.reshape(-1, ch).T.copy()), sr


def write_wav(path: str, wav: np.ndarray, sr: int):
    """float waveform -> 16-bit PCM `.wav` (what `soundfile.write` makes of a float array, api.py:94)."""
    pcm = np.clip(np.round(np.asarray(wav, dtype=np.float64) * 32768.0), -32768, 32767).astype("<i2")
    with wave.open(path, "wb") as w:
        w.setnchannels(1)
        w.setsampwidth(2)
        w.setframerate(sr)
        w.writeframes(pcm.tobytes())


def infer_process(ref_audio, ref_text, gen_text, model_obj, vocoder, mel_spec_type="vocos", show_info=print,
                  progress=None, target_rms=0.1, cross_fade_duration=0.15, nfe_step=32, cfg_strength=2.0,
                  sway_sampling_coef=-1, speed=1.0, fix_duration=None, device=None, tokenize=None):
    """utils_infer.py:453-500: chunk budget from the prompt's bytes per second and the 22 s window, then the batch."""
    audio, sr = read_wav(ref_audio) if isinstance(ref_audio, str) else ref_audio
    secs = audio.shape[-1] / sr
    max_chars = int(len(ref_text.encode("utf-8")) / secs * (22 - secs) * speed)
    gen_text_batches = chunk_text(gen_text, max_chars=max_chars)
    show_info(f"Generating audio in {len(gen_text_batches)} batches...")
    return next(infer_batch_process((audio, sr), ref_text, gen_text_batches, model_obj, vocoder,
                                    mel_spec_type=mel_spec_type, progress=progress, target_rms=target_rms,
                                    cross_fade_duration=cross_fade_duration, nfe_step=nfe_step,
                                    cfg_strength=cfg_strength, sway_sampling_coef=sway_sampling_coef, speed=speed,
                                    fix_duration=fix_duration, device=device, tokenize=tokenize))


class F5TTS:
    def __init__(self, model="F5TTS_v1_Base", ckpt_file="", vocab_file="", ode_method="euler", use_ema=True,
                 vocoder_local_path=None, device=None, hf_cache_dir=None, *, tokenize=None, tokenizer="custom",
                 use_skip_tc=False, use_n2gk_plus=True, allow_uninitialized_vocoder=False):
        backbone, arch, self.mel_spec_type, self.target_sample_rate = model_config(model)
        for k in ("checkpoint_activations",):
            arch.pop(k, None)
        self.ode_method, self.use_ema = ode_method, use_ema
        self.device = device if device is not None else "cuda"
        self.tokenize = tokenize  # list[str] -> list[list[str]]: the reference's convert_char_to_* (host-side Python)
        if not ckpt_file:
            raise ValueError("ckpt_file is required: the reference's hub download (api.py:76-79) needs a network")
        self.vocoder = load_vocoder(self.mel_spec_type, vocoder_local_path is not None, vocoder_local_path or "",
                                    self.device, hf_cache_dir, allow_uninitialized=allow_uninitialized_vocoder)
        self.ema_model = load_model(getattr(backbones, backbone), arch, ckpt_file, self.mel_spec_type, vocab_file,
                                    self.ode_method, self.use_ema, self.device, use_skip_tc=use_skip_tc,
                                    use_n2gk_plus=use_n2gk_plus, tokenizer=tokenizer)

    def transcribe(self, ref_audio, language=None):
        raise NotImplementedError("ASR (whisper) is outside the sampling path; pass ref_text")

    def export_wav(self, wav, file_wave, remove_silence=False):
        if remove_silence:
            raise NotImplementedError("remove_silence needs pydub (outside the sampling path)")
        write_wav(file_wave, wav, self.target_sample_rate)

    def export_spectrogram(self, spec, file_spec):
        np.save(file_spec, spec)  # the reference renders a matplotlib image (utils_infer.py:800-806); raw array here

    def infer(self, ref_file, ref_text, gen_text, show_info=print, progress=None, target_rms=0.1,
              cross_fade_duration=0.15, sway_sampling_coef=-1, cfg_strength=2, nfe_step=32, speed=1.0,
              fix_duration=None, remove_silence=False, file_wave=None, file_spec=None, seed=None):
        if seed is None:
            seed = random.randint(0, sys.maxsize)
        seed_everything(seed)
        self.seed = seed
        if not ref_text:
            raise ValueError("ref_text is required (no ASR here)")
        wav, sr, spec = infer_process(ref_file, ref_text, gen_text, self.ema_model, self.vocoder, self.mel_spec_type,
                                      show_info=show_info, progress=progress, target_rms=target_rms,
                                      cross_fade_duration=cross_fade_duration, nfe_step=nfe_step,
                                      cfg_strength=cfg_strength, sway_sampling_coef=sway_sampling_coef, speed=speed,
                                      fix_duration=fix_duration, device=self.device, tokenize=self.tokenize)
        if file_wave is not None:
            self.export_wav(wav, file_wave, remove_silence)
        if file_spec is not None:
            self.export_spectrogram(spec, file_spec)
        return wav, sr, spec
