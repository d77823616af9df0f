"""`f5b200.api.F5TTS` (reference src/f5_tts/api.py:23-149) and its host helpers.  CPU: the text chunker against golden
vectors produced by the reference's own function (oracle/make_golden_text.py), wav I/O, config table.  GPU: one
`F5TTS.infer` call from files (checkpoint, vocab, 16 kHz stereo prompt) against the oracle's chunk-by-chunk request."""
import json
import os

import numpy as np
import pytest
import torch

import f5b200
from f5b200 import api as A
from helpers import O, max_abs

VOCAB = list(" abcdefghijklmnopqrstuvwxyz,.?!")


def test_chunk_text_matches_reference_golden(golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "chunk_text.json"), encoding="utf-8"))
    assert len(cases) >= 8
    for c in cases:
        assert A.chunk_text(c["text"], c["max_chars"]) == c["chunks"], c["text"]


def test_wav_roundtrip(tmp_path):
    x = np.sin(np.arange(4800) * 0.05) * 0.5
    p = str(tmp_path / "x.wav")
    A.write_wav(p, x, 24000)
    y, sr = A.read_wav(p)
    assert sr == 24000 and y.shape == (1, 4800) and float(np.abs(y[0].numpy() - x).max()) <= 1.0 / 32768 + 1e-7


def test_model_config_table_and_yaml(tmp_path):
    b, arch, mel, sr = A.model_config("F5TTS_v1_Base")
    assert (b, arch["dim"], arch["depth"], arch["heads"], arch["text_mask_padding"], mel, sr) == ("DiT", 1024, 22, 16, True, "vocos", 24000)
    assert A.model_config("E2TTS_Base")[0] == "UNetT" and A.model_config("F5TTS_Small")[1]["dim"] == 768
    y = tmp_path / "tiny.yaml"
    y.write_text("model:\n  backbone: DiT\n  arch: {dim: 128, depth: 2}\n  mel_spec: {mel_spec_type: vocos, target_sample_rate: 24000}\n")
    assert A.model_config(str(y))[1] == {"dim": 128, "depth": 2}
    with pytest.raises(ValueError):
        A.F5TTS(model="F5TTS_v1_Base", ckpt_file="")


@pytest.mark.gpu
def test_f5tts_infer_from_files_vs_oracle(tmp_path):
    from safetensors.torch import save_file

    arch = dict(dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2, text_mask_padding=True, attn_mask_enabled=True)
    cfg = O.ArchCfg(**arch, text_num_embeds=len(VOCAB) + 1)
    sd = O.synth_backbone_weights(cfg, seed=13)
    vsd = O.synth_vocos_weights(5)
    (tmp_path / "vocab.txt").write_text("".join(c + "\n" for c in VOCAB), encoding="utf-8")
    (tmp_path / "tiny.yaml").write_text("model:\n  backbone: DiT\n  arch: " + json.dumps(arch) +
                                        "\n  mel_spec: {mel_spec_type: vocos, target_sample_rate: 24000}\n")
    save_file({"ema_model.transformer." + k: v.contiguous() for k, v in sd.items()}, str(tmp_path / "model.safetensors"))
    g = torch.Generator().manual_seed(3)
    prompt = torch.randn(2, 24000, generator=g) * 0.02  # 1.5 s of 16 kHz stereo
    pcm = np.clip(np.round(prompt.numpy().T * 32768.0), -32768, 32767).astype("<i2")
    import wave
    with wave.open(str(tmp_path / "ref.wav"), "wb") as w:
        w.setnchannels(2); w.setsampwidth(2); w.setframerate(16000); w.writeframes(pcm.tobytes())

    torch.save(vsd, str(tmp_path / "pytorch_model.bin"))  # the released Vocos checkpoint's file name
    with pytest.raises(FileNotFoundError):  # no weights and no network: refuse to vocode with random parameters
        f5b200.load_vocoder("vocos", is_local=False, local_path="", device="cuda")
    tts = A.F5TTS(model=str(tmp_path / "tiny.yaml"), ckpt_file=str(tmp_path / "model.safetensors"),
                  vocab_file=str(tmp_path / "vocab.txt"), device="cuda", vocoder_local_path=str(tmp_path),
                  tokenize=f5b200.tokenizers.char_tokens)  # plain character vocabulary: explicit opt-in
    assert tts.vocoder.precision == "bf16"  # 16-bit kernels by default on CUDA, like the backbone
    tts.ema_model.transformer.float()
    tts.vocoder.precision = "fp32"
    ref_text = "a bc."
    gen_text = "first sentence, quite short. second one is a little bit longer, is it not? third!"
    wav, sr, spec = tts.infer(str(tmp_path / "ref.wav"), ref_text, gen_text, nfe_step=4, seed=11,
                              file_wave=str(tmp_path / "out.wav"), show_info=lambda *_: None)
    assert sr == 24000 and os.path.getsize(tmp_path / "out.wav") == 44 + 2 * len(wav)

    audio16, _ = A.read_wav(str(tmp_path / "ref.wav"))
    audio, rms = O.prepare_ref_audio(audio16, 16000)
    secs = audio16.shape[-1] / 16000
    chunks = A.chunk_text(gen_text, int(len(ref_text.encode()) / secs * (22 - secs)))
    assert len(chunks) >= 2
    ref_len = audio.shape[-1] // 256
    durs = [max(O.request_duration(ref_len, ref_text + " ", c), len(ref_text + " " + c) + 1, ref_len + 2) for c in chunks]
    f5b200.seed_everything(11)
    noises = [torch.randn(d, 100, device="cuda").cpu() for d in durs]
    # the oracle's request starts from the already prepared 24 kHz prompt; undo its own lift with target_rms = rms
    lifted = audio
    want_wave, want_spec = O.infer_request(sd, cfg, vsd, lifted, ref_text, chunks, {c: i for i, c in enumerate(VOCAB)},
                                           noises, steps=4, target_rms=0.0)
    want_wave = want_wave * (rms / 0.1 if rms < 0.1 else 1.0)
    e_mel, e_wav = max_abs(spec, want_spec), max_abs(wav, want_wave)
    print(f"F5TTS.infer from files: {len(chunks)} chunks, mel max-abs {e_mel:.2e}, wave max-abs {e_wav:.2e}")
    assert e_mel <= 1e-3 and e_wav <= 1e-3
