"""the serving surface.  CPU: the socket protocol (float32 PCM pieces, b"END", session re-arm,
first-package rule) with a stand-in synthesiser.  GPU: streaming == non-streaming audio, pieces in text order; the
Triton-contract batch entry against the oracle run request by request."""
import os
import socket
import threading

import numpy as np
import pytest
import torch

import f5b200
from f5b200 import serving as S
from helpers import O, build_cfm, max_abs

VOCAB = {c: i for i, c in enumerate(" abcdefghijklmnopqrstuvwxyz,.?!")}
REF_SOCKET_SERVER = "/root/reference/src/f5_tts/socket_server.py"


def _fake_infer(ref_audio, ref_text, gen_text_batches, model_obj, vocoder, **kw):
    """Deterministic stand-in: chunk k of the request becomes 5000 + 100 k samples of value (k + 1) / 10."""
    assert kw["streaming"] and kw["chunk_size"] == 2048
    for k, _ in enumerate(gen_text_batches):
        w = np.full(5000 + 100 * k, (k + 1) / 10.0, dtype=np.float32)
        for j in range(0, len(w), kw["chunk_size"]):
            yield w[j:j + kw["chunk_size"]], 24000


def _processor(**kw):
    audio = torch.zeros(1, 48000)  # 2 s at 24 kHz
    return S.TTSStreamingProcessor(None, None, (audio, 24000), "twenty bytes of text.", infer_fn=_fake_infer, **kw)


def test_socket_protocol_end_to_end():
    proc = _processor()
    ready, stop, bound = threading.Event(), threading.Event(), []
    th = threading.Thread(target=S.start_server, args=("127.0.0.1", 0, proc, ready, stop, bound), daemon=True)
    th.start()
    assert ready.wait(5)
    try:
        text = "One sentence here. " * 30  # several chunks under the prompt's character budget
        n_chunks = len(_processor().text_batches(text))
        assert n_chunks >= 3
        wav, first, total = S.stream_client("127.0.0.1", bound[0], text)
        want = np.concatenate([np.full(5000 + 100 * k, (k + 1) / 10.0, dtype=np.float32) for k in range(n_chunks)])
        assert wav.dtype == np.float32 and np.array_equal(wav, want)
        assert first is not None and first <= total
        # the raw wire format: float32 little endian pieces, then b"END"; a second text on the same session is NOT re-cut
        with socket.create_connection(("127.0.0.1", bound[0]), timeout=10) as c:
            c.sendall(b"short text.")
            buf = b""
            while not buf.endswith(b"END"):
                buf += c.recv(65536)
            assert (len(buf) - 3) % 4 == 0 and np.frombuffer(buf[:-3], "<f4")[0] == np.float32(0.1)
            assert proc.first_package is False
        for _ in range(50):  # closing the session re-arms the short first package (socket_server.py:189-191)
            if proc.first_package:
                break
            threading.Event().wait(0.05)
        assert proc.first_package is True
    finally:
        stop.set()
        th.join(5)


def test_first_package_rule():
    p = _processor()
    text = "Alpha beta gamma delta. Epsilon zeta eta theta, iota kappa lambda. Mu nu xi omicron! " * 6
    first = p.text_batches(text)
    later = p.text_batches(text)  # same session: plain chunking
    assert later == f5b200.chunk_text(text, max_chars=p.max_chars)
    assert len(first) > len(later) and "".join(first).replace(" ", "") == "".join(later).replace(" ", "")
    assert len(first[0].encode()) <= p.min_chars + 1 or len(first[0].split()) == 1
    assert (p.max_chars, p.few_chars, p.min_chars) == (241, 120, 60)  # 21 bytes / 2 s * 23 s, halves and quarters


@pytest.mark.skipif(not os.path.exists(REF_SOCKET_SERVER), reason="/root/reference is only present in the build container")
def test_first_package_rule_matches_reference_statements():
    """The reference's own statements (socket_server.py:113-121,139-143), lifted with ast and run next to ours."""
    import ast
    import types

    tree = ast.parse(open(REF_SOCKET_SERVER, encoding="utf-8").read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "TTSStreamingProcessor")
    gen = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "generate_stream")
    upd = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "update_reference")
    head = [s for s in gen.body[:2]]  # text_batches = chunk_text(...); if self.first_package: ...
    budget = [s for s in upd.body if isinstance(s, ast.Assign) and not any(isinstance(t, ast.Tuple) for t in s.targets)
              and "torchaudio" not in ast.unparse(s)]
    for secs, ref_text, text in [(2.0, "twenty bytes of text.", "Alpha beta. Gamma delta, epsilon! Zeta eta theta? " * 9),
                                 (3.7, "안녕하세요. 반갑습니다.", "오늘은 날씨가 좋네요. 산책을 갑시다! 어디로 갈까요? " * 7)]:
        audio = torch.zeros(1, int(secs * 24000))
        ours = S.TTSStreamingProcessor(None, None, (audio, 24000), ref_text, infer_fn=_fake_infer)
        me = types.SimpleNamespace(ref_text=ref_text, audio=audio, sr=24000, first_package=True)
        ns = {"self": me, "chunk_text": f5b200.chunk_text, "text": text}
        exec(compile(ast.Module(body=budget, type_ignores=[]), REF_SOCKET_SERVER, "exec"), ns)
        assert (me.max_chars, me.few_chars, me.min_chars) == (ours.max_chars, ours.few_chars, ours.min_chars)
        for _ in range(2):  # first package, then a later one
            exec(compile(ast.Module(body=head, type_ignores=[]), REF_SOCKET_SERVER, "exec"), ns)
            assert ns["text_batches"] == ours.text_batches(text)


def _tiny(mask=True, seed=21):
    cfg = O.ArchCfg(dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2, text_mask_padding=True,
                    attn_mask_enabled=mask, text_num_embeds=len(VOCAB))
    sd = O.synth_backbone_weights(cfg, seed=seed)
    vsd = O.synth_vocos_weights(4)
    cfm = build_cfm(cfg, sd, "fp32")
    cfm.vocab_char_map = VOCAB
    voc = f5b200.Vocos(precision="fp32")
    voc.load_state_dict(vsd)
    return cfg, sd, vsd, cfm, voc.cuda()


@pytest.mark.gpu
def test_streaming_pieces_are_the_request_in_text_order():
    """streaming=True (first chunk alone, later chunks in consecutive batches, two batches in flight) sends exactly the
    audio of the non-streaming request joined without cross-fade, in text order, in pieces of <= 2048 samples."""
    from f5b200 import infer as R

    _, _, _, cfm, voc = _tiny()
    torch.manual_seed(3)
    audio = torch.randn(1, 24000) * 0.03
    ref_text = "the quick brown fox."
    gens = ["jumps over the lazy dog.", "ok?", "and then it sleeps for a while, does it not?", "again and again.", "done!"]
    kw = dict(nfe_step=4, cfg_strength=2.0, sway_sampling_coef=-1.0, tokenize=f5b200.tokenizers.char_tokens)
    torch.manual_seed(9)
    (whole, sr, _), = list(R.infer_batch_process((audio, 24000), ref_text, gens, cfm, voc, cross_fade_duration=0.0, **kw))
    torch.manual_seed(9)
    pieces = list(R.infer_batch_process((audio, 24000), ref_text, gens, cfm, voc, streaming=True, chunk_size=2048, **kw))
    assert all(sr_ == 24000 and 0 < len(p) <= 2048 for p, sr_ in pieces)
    streamed = np.concatenate([p for p, _ in pieces])
    assert streamed.shape == whole.shape
    assert max_abs(streamed, whole) <= 1e-4 * max(1.0, float(np.abs(whole).max()))


@pytest.mark.gpu
def test_socket_server_with_the_real_path():
    _, _, _, cfm, voc = _tiny()
    audio = torch.randn(1, 36000, generator=torch.Generator().manual_seed(1)) * 0.05
    proc = S.TTSStreamingProcessor(cfm, voc, (audio, 24000), "a quick prompt, yes.", device="cuda", nfe_step=4,
                                   tokenize=f5b200.tokenizers.char_tokens)
    ready, stop, bound = threading.Event(), threading.Event(), []
    th = threading.Thread(target=S.start_server, args=("127.0.0.1", 0, proc, ready, stop, bound), daemon=True)
    th.start()
    assert ready.wait(5)
    try:
        wav, first, total = S.stream_client("127.0.0.1", bound[0], "hello there. how are you today? fine, thanks. and you?")
        assert len(wav) > 24000 and np.isfinite(wav).all() and first < total
        print(f"socket stream: {len(wav) / 24000:.2f} s of audio, first bytes after {first * 1e3:.1f} ms, END after {total * 1e3:.1f} ms")
    finally:
        stop.set()
        th.join(5)


@pytest.mark.gpu
def test_triton_contract_batch_vs_oracle_one_by_one():
    """execute([req0, req1]) -- different prompts, one variable-length batch -- equals the oracle run request by request
    (prompt prepared, mel, sample with the same noise rows, Vocos, rms rescale)."""
    cfg, sd, vsd, cfm, voc = _tiny(seed=23)
    g = torch.Generator().manual_seed(6)
    reqs = []
    for n, scale, ref, tgt in [(30000, 0.02, "first prompt text.", "say this sentence, please."),
                               (41000, 0.3, "another, longer prompt here.", "short one.")]:
        wav = (torch.randn(1, n + 500, generator=g) * scale).numpy()
        reqs.append(dict(reference_wav=wav, reference_wav_len=np.array([[n]], dtype=np.int32),
                         reference_text=np.array([[ref.encode()]], dtype=object), target_text=tgt))
    model = S.TritonLikeModel(cfm, voc, device="cuda", tokenize=f5b200.tokenizers.char_tokens, nfe_step=4)
    torch.manual_seed(12)
    got = model.execute(reqs)
    # oracle, one request at a time, same noise stream (rows are drawn in batch order on the device)
    texts = [S.TritonLikeModel._text(r["reference_text"]) for r in reqs], [r["target_text"] for r in reqs]
    prep = []
    for r in reqs:
        n = int(r["reference_wav_len"][0, 0])
        audio, rms = O.prepare_ref_audio(torch.from_numpy(r["reference_wav"])[:, :n], 24000)
        prep.append((O.mel_spectrogram(audio).permute(0, 2, 1), float(rms)))
    est = [int(m.shape[1] * (1 + len(t.encode()) / len(rt.encode()))) for (m, _), rt, t in zip(prep, *texts)]
    ids = [O.list_str_to_idx([list(rt + t)], VOCAB) for rt, t in zip(*texts)]
    durs = [max(e, max(int(i.shape[1]), m.shape[1]) + 1) for e, i, (m, _) in zip(est, ids, prep)]
    torch.manual_seed(12)
    noises = [torch.randn(d, 100, device="cuda").cpu() for d in durs]
    for k, ((mel, rms), e) in enumerate(zip(prep, est)):
        ref = O.cfm_sample(sd, cfg, mel, ids[k], e, steps=4, cfg_strength=2.0, sway_sampling_coef=-1.0, y0=noises[k][None])
        want = O.vocos_decode(vsd, ref.out[:, mel.shape[1]: e].permute(0, 2, 1))[0].numpy()
        if rms < 0.1:
            want = want * (rms / 0.1)
        assert got[k].shape == want.shape
        err = max_abs(got[k], want)
        print(f"triton contract request {k}: {len(want)} samples, wave max-abs {err:.2e}")
        assert err <= 1e-3 * max(1.0, float(np.abs(want).max()))
