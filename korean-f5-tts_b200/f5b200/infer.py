"""Request level of the path: `infer_batch_process` with the reference's signature and results
(src/f5_tts/infer/utils_infer.py:500-770), but every text chunk of the request goes through ONE variable-length
`CFM.sample` batch (padded masks, per-row durations) instead of a thread pool of batch-1 calls that share the
backbone's text cache (`utils_infer.py:725-727`, `backbones/dit.py:181,259-269`).

The reference-audio preparation (row f2: channel mix-down, RMS lift, resampling to 24 kHz) is one library call,
`f5b_ref_audio`.  What stays host-side Python, exactly as in the reference: the duration rule
(`utils_infer.py:619-626`), tokenisation (a callable handed in by the caller: the reference's `convert_char_to_*`
functions work unchanged), RMS rescale of the result and the linear cross-fade (`utils_infer.py:734-770`).
The mel front end, the ODE and the Vocos decode run in libf5b200.so; nothing here computes them on the CPU."""
from __future__ import annotations

from typing import Callable, Iterable, List, Sequence

import numpy as np
import torch

from . import tokenizers
from .sharding import micro_batches
from .utils import list_str_to_idx, list_str_to_tensor

target_sample_rate = 24000  # utils_infer.py:62-66
hop_length = 256
TARGET_RMS = 0.1
CROSS_FADE_DURATION = 0.15


def chunk_duration(ref_audio_len: int, ref_text: str, gen_text: str, speed: float = 1.0, fix_duration=None) -> int:
    """Frames of prompt + generated speech for one chunk (utils_infer.py:541-545, 619-626): utf-8 byte lengths,
    texts under 10 bytes are slowed to 0.3."""
    if fix_duration is not None:
        return int(fix_duration * target_sample_rate / hop_length)
    local_speed = 0.3 if len(gen_text.encode("utf-8")) < 10 else speed
    ref_text_len = len(ref_text.encode("utf-8"))
    gen_text_len = len(gen_text.encode("utf-8"))
    return ref_audio_len + int(ref_audio_len / ref_text_len * gen_text_len / local_speed)


def cross_fade(waves: Sequence[np.ndarray], cross_fade_duration: float = CROSS_FADE_DURATION,
               sample_rate: int = target_sample_rate) -> np.ndarray:
    """Join chunk waveforms with a linear cross-fade (utils_infer.py:734-770); plain concatenation when the duration
    is <= 0 or a chunk is empty."""
    if cross_fade_duration <= 0:
        return np.concatenate(list(waves))
    final = waves[0]
    for nxt in waves[1:]:
        n = min(int(cross_fade_duration * sample_rate), len(final), len(nxt))
        if n <= 0:
            final = np.concatenate([final, nxt])
            continue
        fade_out = np.linspace(1, 0, n)
        fade_in = np.linspace(0, 1, n)
        final = np.concatenate([final[:-n], final[-n:] * fade_out + nxt[:n] * fade_in, nxt[n:]])
    return final


def sinc_filter_bank(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """The filter bank of `torchaudio.transforms.Resample(orig, new)` with its defaults (Hann-windowed sinc, width 6,
    roll-off 0.99; third-party torchaudio `functional._get_sinc_resample_kernel`, the reference's call is
    utils_infer.py:530-532): float64 construction, float32 result [n_new, 2 * width + n_orig].  Host-side, cached."""
    import math

    g = math.gcd(int(orig_freq), int(new_freq))
    o, n = int(orig_freq) // g, int(new_freq) // g
    base_freq = min(o, n) * rolloff
    width = math.ceil(lowpass_filter_width * o / base_freq)
    idx = torch.arange(-width, width + o, dtype=torch.float64)[None, None] / o
    # (the phase offsets pass through float32 there: integer arange / int, then promoted by the float64 `idx`)
    t = (torch.arange(0, -n, -1)[:, None, None] / n).to(torch.float32) + idx
    t = (t * base_freq).clamp_(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    scale = base_freq / o
    bank = torch.where(t == 0, torch.tensor(1.0, dtype=torch.float64), t.sin() / t) * window * scale
    return bank.to(torch.float32).reshape(n, 2 * width + o).contiguous(), o, n, width


_BANKS: dict = {}


def prepare_ref_audio(ref_audio, target_rms: float = TARGET_RMS, device=None):
    """Mono mix, RMS lift, resample to 24 kHz (utils_infer.py:523-533) in ONE library call (`f5b_ref_audio`): the
    waveform crosses to the device once, at its original rate.  Returns (audio [1, nw24] float32 on `device`, rms)."""
    import ctypes as C

    from . import _lib

    audio, sr = ref_audio
    if audio.ndim == 1:
        audio = audio[None]
    dev = torch.device(device if device is not None else (audio.device if audio.is_cuda else "cuda"))
    _lib.device_index(dev)  # no CPU fallback: a non-CUDA device is an error
    wav = audio.to(torch.float32).to(dev, non_blocking=True).contiguous()
    ch, nw = wav.shape
    key = (int(sr), str(dev))
    if key not in _BANKS:
        bank, o, n, width = (sinc_filter_bank(sr, target_sample_rate) if sr != target_sample_rate
                             else (torch.ones(1, 1), 1, 1, 0))
        _BANKS[key] = (bank.to(dev), o, n, width)
    bank, o, n, width = _BANKS[key]
    out_len = -(-n * nw // o)  # ceil(new * length / orig), torchaudio _apply_sinc_resample_kernel
    out = torch.empty((1, out_len), device=dev, dtype=torch.float32)
    rms = C.c_float(0.0)
    _lib.check(_lib.lib().f5b_ref_audio(wav.data_ptr(), ch, nw, bank.data_ptr(), n, bank.shape[1], o, width,
                                        float(target_rms), out.data_ptr(), out_len, C.byref(rms),
                                        _lib.device_index(dev), _lib.MEM_DEVICE, _lib.current_stream_ptr(dev)))
    return out, float(rms.value)


def infer_batch_process(ref_audio, ref_text: str, gen_text_batches: Iterable[str], model_obj, vocoder,
                        mel_spec_type="vocos", progress=None, target_rms=TARGET_RMS,
                        cross_fade_duration=CROSS_FADE_DURATION, nfe_step=32, cfg_strength=2.0, sway_sampling_coef=-1,
                        speed=1, fix_duration=None, device=None, streaming=False, chunk_size=2048, *,
                        tokenize: Callable[[List[str]], List[List[str]]] | None = None, max_rows: int = 16384,
                        max_batch: int = 64):
    """Generator with the reference's contract: yields `(final_wave, 24000, combined_spectrogram)` once (or
    `(chunk, 24000)` pieces when `streaming`).  Keyword-only extras: `tokenize` (list[str] -> list[list[str]]; by default the
    reference's `convert_char_to_*` picked by the reference's rule from `model_obj._tokenizer_type / _use_skip_tc /
    _use_n2gk_plus` and the vocabulary, `f5b200.tokenizers`), `max_rows` / `max_batch` (micro-batch limits in padded
    frames / utterances).

    Chunks are grouped into length-sorted micro-batches and each group is ONE `sample()`; rows of different length
    need the backbone's key mask (`attn_mask_enabled=True`) to equal their batch-1 results, so a backbone built without
    it gets one chunk per `sample()` call (same kernels, batch 1).  Noise rows are drawn chunk by chunk from the
    global generator in request order, i.e. the stream of `torch.randn` calls is that of the reference's sequential
    (streaming) loop."""
    if mel_spec_type != "vocos":
        raise ValueError("only the Vocos mel / vocoder pair is built (BigVGAN is out of scope)")
    gen_text_batches = list(gen_text_batches)
    device = device if device is not None else model_obj.device
    if not gen_text_batches:
        yield None, target_sample_rate, None
        return
    audio, rms = prepare_ref_audio(ref_audio, target_rms, device)
    if len(ref_text[-1].encode("utf-8")) == 1:
        ref_text = ref_text + " "
    if tokenize is None:  # the reference's rule (utils_infer.py:549-676); raises when its converters are not importable
        tokenize = tokenizers.resolve(tokenizers.choose_for_model(model_obj))

    ref_audio_len = audio.shape[-1] // hop_length
    durations = [chunk_duration(ref_audio_len, ref_text, g, speed, fix_duration) for g in gen_text_batches]
    tokens = tokenize([ref_text + g for g in gen_text_batches])
    vocab = getattr(model_obj, "vocab_char_map", None)
    ids = [(list_str_to_idx([t], vocab) if vocab is not None else list_str_to_tensor([t]))[0] for t in tokens]

    # prompt mel once per request (the reference recomputes it inside every sample() call, cfm.py:106-109)
    ref_mel = model_obj.mel_spec(audio).permute(0, 2, 1).to(torch.float32)  # [1, T, 100]
    n_ch = ref_mel.shape[-1]
    # the effective length of a row after cfm.py:135-137
    eff = [max(max(int(len(i)), ref_mel.shape[1]) + 1, d) for i, d in zip(ids, durations)]

    # noise in request order (cfm.py:196-201 with seed None)
    noise = [torch.randn(e, n_ch, device=device, dtype=torch.float32) for e in eff]

    ragged_ok = bool(getattr(model_obj.transformer, "attn_mask_enabled", False))
    order = list(range(len(eff)))
    if streaming:
        # audio must leave in text order and the FIRST piece as early as possible (socket_server.py:138-177): the first
        # chunk runs alone, the rest in consecutive (unsorted) runs, each run one sample() + one Vocos call
        groups, cur = [[0]], []
        for i in order[1:]:
            can_join = cur and (ragged_ok or eff[i] == eff[cur[0]]) and len(cur) < max_batch and \
                (len(cur) + 1) * max(eff[i], max(eff[k] for k in cur)) <= max_rows
            if cur and not can_join:
                groups.append(cur)
                cur = []
            cur.append(i)
        if cur:
            groups.append(cur)
    elif ragged_ok:
        groups = micro_batches(order, eff, max_rows=max_rows, max_batch=max_batch)
    else:  # equal-length rows may still share a batch
        by_len: dict = {}
        for i in order:
            by_len.setdefault(eff[i], []).append(i)
        groups = [g[k:k + max_batch] for g in by_len.values() for k in range(0, len(g), max_batch)]

    # every shape of the request is known here: reserve the workspaces once, so that no micro-batch re-allocates (a
    # re-allocation synchronises the device and drops the cached graphs) -- objects without the method are left alone
    if len(groups) > 1:
        tr = getattr(model_obj, "transformer", None)
        if hasattr(tr, "reserve"):
            tr.reserve([(len(g), max(eff[i] for i in g), max(len(ids[i]) for i in g)) for g in groups], steps=nfe_step)
        if hasattr(vocoder, "reserve"):
            for g in groups:
                vocoder.reserve(len(g), max(max(eff[i] for i in g) - ref_audio_len, 2))

    def launch(g):
        """Enqueue one micro-batch: sample() + ONE Vocos call over its rows (per-row lengths: each row is decoded as if
        alone) + device -> pinned-host copies.  Nothing here waits for the GPU (the C ABI is asynchronous for device
        buffers); the returned event marks the copies' end."""
        b, n = len(g), max(eff[i] for i in g)
        nt = max(len(ids[i]) for i in g)
        text = torch.full((b, nt), -1, dtype=torch.long)
        y0 = torch.zeros((b, n, n_ch), device=device, dtype=torch.float32)
        for r, i in enumerate(g):
            text[r, :len(ids[i])] = ids[i]
            y0[r, :eff[i]] = noise[i]
        dur = torch.tensor([durations[i] for i in g], device=device, dtype=torch.long)
        glen = torch.tensor([eff[i] - ref_audio_len for i in g], dtype=torch.int32)
        with torch.inference_mode():
            out, _ = model_obj.sample(cond=ref_mel.expand(b, -1, -1).contiguous(), text=text.to(device),
                                      duration=dur if b > 1 else int(dur[0]), steps=nfe_step,
                                      cfg_strength=cfg_strength, sway_sampling_coef=sway_sampling_coef, y0=y0,
                                      return_trajectory=False)
            gen = out[:, ref_audio_len:, :].to(torch.float32).permute(0, 2, 1).contiguous()  # [b, 100, n - ref]
            wav = vocoder.decode(gen, glen.to(device)) if b > 1 else vocoder.decode(gen)
            if rms < target_rms:
                wav = wav * rms / target_rms
            wav_h = torch.empty(wav.shape, dtype=wav.dtype, pin_memory=True)
            gen_h = torch.empty(gen.shape, dtype=gen.dtype, pin_memory=True)
            wav_h.copy_(wav, non_blocking=True)
            gen_h.copy_(gen, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        return g, glen, wav_h, gen_h, ev

    def collect(handle):
        g, glen, wav_h, gen_h, ev = handle
        ev.synchronize()
        wav_n, gen_n = wav_h.numpy(), gen_h.numpy()
        return [(i, wav_n[r, : (int(glen[r]) - 1) * hop_length], gen_n[r, :, : int(glen[r])]) for r, i in enumerate(g)]

    # A worker thread enqueues the micro-batches one after the other (at most two ahead of the consumer); this thread
    # waits for each batch's copies and hands the audio out.  The launches need their own thread: without CUDA graphs a
    # sample() is ~6000 kernel launches and the driver's launch queue holds ~1000, so the enqueueing call itself only
    # returns when the batch is nearly done -- measured: the first chunk of a streaming request left after 176 ms (the
    # whole request) instead of 42 ms while the second batch was being enqueued from the consumer's thread.
    import queue
    import threading

    mels: List[np.ndarray | None] = [None] * len(eff)
    waves: List[np.ndarray | None] = [None] * len(eff)
    handles: "queue.Queue" = queue.Queue(maxsize=2)
    stop = threading.Event()

    parent_stream = torch.cuda.current_stream(torch.device(device))  # current device / stream are per thread

    def producer():
        try:
            with torch.cuda.device(parent_stream.device), torch.cuda.stream(parent_stream):
                for g in (progress.tqdm(groups) if progress is not None else groups):
                    if stop.is_set():
                        break
                    handles.put(launch(g))
            handles.put(None)
        except BaseException as e:  # noqa: BLE001  (re-raised in the consumer)
            handles.put(e)

    worker = threading.Thread(target=producer, daemon=True)
    worker.start()
    try:
        while True:
            h = handles.get()
            if h is None:
                break
            if isinstance(h, BaseException):
                raise h
            for i, w, m in collect(h):
                waves[i], mels[i] = w, m
                if streaming:  # (groups are consecutive and in order)
                    for j in range(0, len(w), chunk_size):
                        yield w[j:j + chunk_size], target_sample_rate
    finally:
        stop.set()
        while worker.is_alive():  # unblock a producer waiting on the full queue (generator closed early)
            try:
                handles.get_nowait()
            except queue.Empty:
                worker.join(0.01)
    if streaming:
        return
    yield cross_fade(waves, cross_fade_duration), target_sample_rate, np.concatenate(mels, axis=1)
