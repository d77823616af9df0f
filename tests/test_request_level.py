"""the request level (`infer_batch_process`) -- all text chunks of a request as one
variable-length batch.  CPU: host logic against the oracle's restatement of utils_infer.py.  GPU: the batched request
against the oracle running the reference's chunk-by-chunk loop on the same audio, text and noise."""
import numpy as np
import pytest
import torch

import f5b200
from f5b200 import infer as R
from f5b200 import tokenizers as TK
from helpers import O, build_cfm, max_abs

VOCAB = {c: i for i, c in enumerate(" abcdefghijklmnopqrstuvwxyz,.?!")}


@pytest.mark.parametrize("ref_len,ref,gen,speed,fix", [(282, "some call me nature. ", "hello there, how are you today?", 1.0, None),
                                                       (282, "안녕하세요. ", "짧다", 1.0, None),  # < 10 bytes -> 0.3
                                                       (100, "abc. ", "a much longer sentence than the prompt was.", 0.7, None),
                                                       (100, "abc. ", "ignored", 1.0, 4.0)])
def test_duration_rule(ref_len, ref, gen, speed, fix):
    assert R.chunk_duration(ref_len, ref, gen, speed, fix) == O.request_duration(ref_len, ref, gen, speed, fix)


@pytest.mark.parametrize("dur", [0.15, 0.0, 10.0])
def test_cross_fade_matches_oracle(dur):
    rng = np.random.default_rng(0)
    waves = [rng.standard_normal(n).astype(np.float32) for n in (9000, 2000, 5000, 0 if dur == 10.0 else 3000)]
    a, b = R.cross_fade(waves, dur), O.linear_cross_fade(waves, dur)
    assert a.shape == b.shape and np.array_equal(a, b)
    if dur == 0.15:
        assert len(a) == sum(map(len, waves)) - (2000 + 3600 + 3000)  # overlap clipped to the shorter chunk


def _request_golden(golden_dir):
    import json
    import os
    with open(os.path.join(golden_dir, "request_level.json"), encoding="utf-8") as f:
        return json.load(f)


def test_duration_rule_matches_reference_golden(golden_dir):
    """Pinned to the reference's own statements (utils_infer.py:541-545,678-685 lifted by oracle/make_golden_request.py)."""
    for c in _request_golden(golden_dir)["duration"]:
        ref_len = c["n_audio"] // 256
        assert ref_len == c["ref_audio_len"]
        for fn in (R.chunk_duration, O.request_duration):
            assert fn(ref_len, c["ref_text"], c["gen_text"], c["speed"], c["fix_duration"]) == c["duration"], c


def test_cross_fade_matches_reference_golden(golden_dir):
    """The reference's cross-fade block (utils_infer.py:734-770, lifted) on the same seeded waves, bit for bit."""
    rng = np.random.default_rng(0)
    for c in _request_golden(golden_dir)["cross_fade"]:
        waves = [rng.standard_normal(n).astype(np.float32) for n in c["lens"]]
        want = np.asarray(c["final"], dtype=np.float64)
        for fn in (R.cross_fade, O.linear_cross_fade):
            got = fn(waves, c["cross_fade_duration"], c["sample_rate"])
            assert got.shape == want.shape and np.array_equal(np.asarray(got, dtype=np.float64), want), c["lens"]


def test_tokenizer_dispatch_matches_reference_golden(golden_dir):
    """Every (tokenizer type, vocabulary flavour, skipTC, N2gk+) combination picks the converter, the keyword arguments
    and the normalisation the reference's dispatch (utils_infer.py:549-676, lifted) picks."""
    from oracle.make_golden_request import VOCABS

    rows = _request_golden(golden_dir)["dispatch"]
    assert len(rows) > 300
    for r in rows:
        vocab = None if r["vocab"] == "none" else VOCABS[r["vocab"]]
        ch = TK.choose(r["tokenizer_type"], vocab, r["skip_tc"], r["n2gk"], r.get("legacy", False))
        want_norm = any(c["fn"] == "normalize_n2gk_plus" for c in r["calls"])
        conv = [c for c in r["calls"] if c["fn"] != "normalize_n2gk_plus"]
        assert len(conv) == 1
        assert (ch.normalize, ch.fn, ch.kwargs) == (want_norm, conv[0]["fn"], conv[0]["kwargs"]), r


def test_tokenizer_is_never_silently_characters():
    """with a Korean vocabulary and no converter importable the request must fail, not emit characters."""
    class _M:
        device = "cpu"
        vocab_char_map = {"k": 0, "kⁱ": 1}
        _tokenizer_type = "custom"
    assert TK.choose_for_model(_M()).fn == "convert_char_to_allophone"
    if TK._reference_utils() is None:
        with pytest.raises(RuntimeError, match="convert_char_to_allophone"):
            TK.resolve(TK.choose_for_model(_M()))
    # a stand-in module shows the binding: normalisation first, then the converter with the named filter constants
    import types
    calls = []
    mod = types.SimpleNamespace(PHONEMES_N=["n"], normalize_n2gk_plus=lambda t: t.upper(),
                                convert_char_to_allophone=lambda tl, **kw: calls.append((tl, kw)) or [list(t) for t in tl])
    fn = TK.resolve(TK.choose("kor_n_only", {"a": 0}, False, True), module=mod)
    assert fn(["ab"]) == [["A", "B"]]
    assert calls == [(["AB"], dict(apply_init=False, apply_pal=False, apply_coda=True, coda_filter=["n"], initial_filter=None))]


@pytest.mark.parametrize("sr", [44100, 22050, 16000, 48000, 8000])
def test_sinc_filter_bank_is_torchaudio_s(sr):
    """The host-built filter bank equals the one torchaudio.transforms.Resample holds (the reference's call)."""
    import torchaudio

    ref = torchaudio.transforms.Resample(sr, 24000)
    bank, o, n, width = R.sinc_filter_bank(sr, 24000)
    assert (o, n, width) == (ref.orig_freq // ref.gcd, ref.new_freq // ref.gcd, ref.width)
    assert torch.equal(bank, ref.kernel.reshape(n, -1))


@pytest.mark.gpu
@pytest.mark.parametrize("sr,ch,nw,scale", [(44100, 2, 132300, 0.02), (24000, 1, 72000, 0.02), (16000, 1, 50001, 0.5),
                                            (48000, 2, 96000, 0.01), (22050, 1, 4097, 0.03), (8000, 3, 777, 0.05)])
def test_ref_audio_front_half_vs_torchaudio(sr, ch, nw, scale):
    """f5b_ref_audio (mix-down + rms lift + polyphase resampler) against the reference's torch / torchaudio calls."""
    g = torch.Generator().manual_seed(nw)
    wav = torch.randn(ch, nw, generator=g) * scale
    want, want_rms = O.prepare_ref_audio(wav, sr)
    got, rms = R.prepare_ref_audio((wav, sr), device="cuda")
    assert got.shape == want.shape
    err = max_abs(got, want)
    print(f"ref audio {sr} Hz x{ch}: rms {rms:.6f} vs {want_rms:.6f}, max-abs {err:.2e} (peak {float(want.abs().max()):.2f})")
    assert abs(rms - want_rms) <= 1e-6 * max(1.0, want_rms) and err <= 2e-6 * max(1.0, float(want.abs().max()))
    again, _ = R.prepare_ref_audio((wav.cuda(), sr))  # device input, cached filter bank: same bits
    assert torch.equal(again, got)


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_ref_audio_has_no_cpu_path():
    from f5b200 import F5BError

    with pytest.raises(F5BError):
        R.prepare_ref_audio((torch.randn(1, 2400), 24000), device="cpu")


def test_empty_request_yields_none():
    class _M:
        device = "cpu"
    assert list(R.infer_batch_process((torch.randn(1, 2400) * 0.01, 24000), "a. ", [], _M(), None)) == [(None, 24000, None)]


@pytest.mark.gpu
@pytest.mark.parametrize("mask_enabled", [True, False])
def test_batched_request_vs_oracle_chunk_loop(mask_enabled):
    import f5b200

    cfg = O.ArchCfg.named("F5TTS_v1_Base", depth=2, attn_mask_enabled=mask_enabled, text_num_embeds=len(VOCAB))
    sd = O.synth_backbone_weights(cfg, seed=11)
    vsd = O.synth_vocos_weights(3)
    torch.manual_seed(5)
    audio = torch.randn(1, 24000) * 0.03  # 1 s prompt, rms < 0.1 -> lifted and rescaled back
    ref_text = "the quick brown fox."
    gens = ["jumps over the lazy dog, twice.", "ok?", "and then it sleeps for a while, does it not?", "jumps over the lazy dog, thrice"]
    cfm = build_cfm(cfg, sd, "fp32")
    cfm.vocab_char_map = VOCAB
    voc = f5b200.Vocos(precision="fp32")
    voc.load_state_dict(vsd)
    voc = voc.cuda()

    torch.manual_seed(21)
    (wave, sr, spec), = list(R.infer_batch_process((audio, 24000), ref_text, gens, cfm, voc, nfe_step=4, cfg_strength=2.0,
                                                   sway_sampling_coef=-1.0, tokenize=f5b200.tokenizers.char_tokens))
    # the same noise stream, chunk by chunk in request order
    ref_len = 24000 // 256
    torch.manual_seed(21)
    durs = [max(O.request_duration(ref_len, ref_text + " ", g), len(ref_text + " " + g) + 1, ref_len + 2) for g in gens]
    noises = [torch.randn(d, 100, device="cuda").cpu() for d in durs]
    want_wave, want_spec = O.infer_request(sd, cfg, vsd, audio, ref_text, gens, VOCAB, noises, steps=4)
    assert sr == 24000 and wave.shape == want_wave.shape and spec.shape == want_spec.shape
    e_mel, e_wav = max_abs(spec, want_spec), max_abs(wave, want_wave)
    print(f"request level (attn mask {mask_enabled}): mel max-abs {e_mel:.3e}, wave max-abs {e_wav:.3e}")
    assert e_mel <= 1e-3 and e_wav <= 1e-3
