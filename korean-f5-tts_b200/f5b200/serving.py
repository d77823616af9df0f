"""Serving surface of the path: the reference's two deployment front ends on top of the batched
request level (`infer.infer_batch_process`).

1. The socket streaming server (reference src/f5_tts/socket_server.py:70-214).  Same wire protocol: the client sends
   utf-8 text (one `recv(1024)`), the server answers with raw little-endian float32 PCM at 24 kHz in pieces of at most
   2048 samples as they become available and finishes the utterance with the 3 bytes b"END" (:158-174); an empty read
   closes the session and re-arms the short first package (:189-191).  Same first-package rule: the first text chunk of
   a session is re-cut with half and a quarter of the character budget so the first audio leaves early (:139-143).
   What differs is underneath: the first chunk runs alone, all later chunks run as variable-length batches while the
   earlier audio is still being sent (`infer_batch_process(streaming=True)`).

2. The Triton python-backend contract (reference runtime/triton_trtllm/model_repo_f5_tts/f5_tts/config.pbtxt:44-74,
   1/model.py:176-268): inputs `reference_wav` fp32 [1, n], `reference_wav_len` int32 [1, 1], `reference_text`,
   `target_text` -> output `waveform` fp32 [m].  `TritonLikeModel.execute` takes the dynamic batcher's request list and
   runs ALL of it as one variable-length `sample()` + one Vocos call (the reference's TRT-LLM engine pads to the batch
   maximum the same way); the length rule is the reference's `ref_len * (1 + len(target) / len(reference))` in utf-8
   bytes (:222-227).  No Triton runtime is needed to call it; a `model.py` shim would forward `execute` one to one.
"""
from __future__ import annotations

import logging
import socket
import threading
import time
from typing import Callable, List, Optional, Sequence

import numpy as np
import torch

from .api import chunk_text, read_wav
from .infer import hop_length, infer_batch_process, prepare_ref_audio, target_sample_rate
from .utils import list_str_to_idx

logger = logging.getLogger(__name__)

END_MARK = b"END"  # socket_server.py:174


class TTSStreamingProcessor:
    """socket_server.py:70-177 with the model / vocoder objects handed in (they are built by `load_model` /
    `load_vocoder`, or by the caller).  `infer_fn` is `infer_batch_process` unless a test injects a stand-in."""

    def __init__(self, model, vocoder, ref_audio, ref_text: str, device=None, tokenize=None, nfe_step: int = 32,
                 cfg_strength: float = 2.0, sway_sampling_coef: float = -1.0, chunk_size: int = 2048,
                 infer_fn: Optional[Callable] = None, warm_up: bool = True):
        self.model, self.vocoder, self.device, self.tokenize = model, vocoder, device, tokenize
        self.nfe_step, self.cfg_strength, self.sway = nfe_step, cfg_strength, sway_sampling_coef
        self.chunk_size = chunk_size
        self.infer_fn = infer_fn if infer_fn is not None else infer_batch_process
        self.sampling_rate = target_sample_rate
        self.first_package = True
        self.last_first_chunk_s: Optional[float] = None  # time from request to the first PCM bytes (last request)
        self.update_reference(ref_audio, ref_text)
        if warm_up:
            self._warm_up()

    def update_reference(self, ref_audio, ref_text: str):
        """:113-121 -- character budgets from the prompt's bytes per second and a 25 s window."""
        self.audio, self.sr = read_wav(ref_audio) if isinstance(ref_audio, str) else ref_audio
        self.ref_text = ref_text
        secs = self.audio.shape[-1] / self.sr
        per_s = len(ref_text.encode("utf-8")) / secs
        self.max_chars = int(per_s * (25 - secs))
        self.few_chars = int(per_s * (25 - secs) / 2)
        self.min_chars = int(per_s * (25 - secs) / 4)

    def _stream(self, text_batches: Sequence[str]):
        return self.infer_fn((self.audio, self.sr), self.ref_text, list(text_batches), self.model, self.vocoder,
                             progress=None, device=self.device, streaming=True, chunk_size=self.chunk_size,
                             nfe_step=self.nfe_step, cfg_strength=self.cfg_strength, sway_sampling_coef=self.sway,
                             tokenize=self.tokenize)

    def _warm_up(self):
        for _ in self._stream(["Warm-up text for the model."]):  # :123-136
            pass

    def text_batches(self, text: str) -> List[str]:
        """:139-143 -- the first package of a session is cut finer so that its first chunk is short."""
        batches = chunk_text(text, max_chars=self.max_chars)
        if self.first_package and batches:
            batches = chunk_text(batches[0], max_chars=self.few_chars) + batches[1:]
            batches = chunk_text(batches[0], max_chars=self.min_chars) + batches[1:]
            self.first_package = False
        return batches

    def generate_stream(self, text: str, conn) -> int:
        """Synthesize `text` and send it down `conn` (anything with `sendall`).  Returns the number of samples sent."""
        t0 = time.perf_counter()
        sent, first = 0, None
        for chunk, _ in self._stream(self.text_batches(text)):
            if len(chunk) > 0:
                # struct.pack(f"{n}f", *chunk) of the reference == the array's native float32 bytes
                conn.sendall(np.ascontiguousarray(chunk, dtype="<f4").tobytes())
                if first is None:
                    first = time.perf_counter() - t0
                sent += len(chunk)
        conn.sendall(END_MARK)
        self.last_first_chunk_s = first
        return sent


def handle_client(conn, processor: TTSStreamingProcessor):
    """:180-204 -- one session: texts until the peer closes; errors end the session, not the server."""
    try:
        with conn:
            try:
                conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            except OSError:
                pass
            while True:
                data = conn.recv(1024)
                if not data:
                    processor.first_package = True
                    break
                text = data.decode("utf-8").strip()
                try:
                    processor.generate_stream(text, conn)
                except Exception:  # noqa: BLE001  (the reference logs and drops the session, :198-201)
                    logger.exception("error during synthesis")
                    break
    except Exception:  # noqa: BLE001
        logger.exception("error handling client")


def start_server(host: str, port: int, processor: TTSStreamingProcessor, ready: Optional[threading.Event] = None,
                 stop: Optional[threading.Event] = None, bound: Optional[list] = None):
    """:207-214.  `ready` / `stop` / `bound` (extras) let a test run the loop in a thread on an ephemeral port."""
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        s.bind((host, port))
        s.listen()
        s.settimeout(0.2)
        if bound is not None:
            bound.append(s.getsockname()[1])
        if ready is not None:
            ready.set()
        while stop is None or not stop.is_set():
            try:
                conn, _ = s.accept()
            except socket.timeout:
                continue
            conn.settimeout(None)
            handle_client(conn, processor)


def stream_client(host: str, port: int, text: str, timeout: float = 120.0):
    """Minimal client of the protocol (what the reference's README client does): returns (float32 waveform,
    seconds until the first bytes, seconds until END)."""
    t0 = time.perf_counter()
    first = None
    buf = bytearray()
    with socket.create_connection((host, port), timeout=timeout) as c:
        c.sendall(text.encode("utf-8"))
        while True:
            data = c.recv(65536)
            if not data:
                break
            if first is None:
                first = time.perf_counter() - t0
            buf += data
            if buf[-3:] == END_MARK and (len(buf) - 3) % 4 == 0:
                # float32 PCM cannot end in b"END" by accident at a 4-byte boundary AND be followed by silence on the
                # socket; the reference's client uses the same test
                break
    total = time.perf_counter() - t0
    return np.frombuffer(bytes(buf[:-3]), dtype="<f4").copy(), first, total


class TritonLikeModel:
    """`TritonPythonModel.execute` of the reference's f5_tts model repository (1/model.py:176-268) over the B200 path.
    A request is a mapping with the config.pbtxt inputs; the result list holds one float32 `waveform` per request."""

    def __init__(self, model, vocoder, device=None, tokenize=None, reference_audio_sample_rate: int = 24000,
                 target_rms: float = 0.1, nfe_step: int = 16, cfg_strength: float = 2.0,
                 sway_sampling_coef: float = -1.0, max_mel_len: int = 4096):
        from . import tokenizers

        self.model, self.vocoder = model, vocoder
        self.device = device if device is not None else model.device
        self.tokenize = tokenize if tokenize is not None else tokenizers.resolve(tokenizers.choose_for_model(model))
        self.sr, self.target_rms = reference_audio_sample_rate, target_rms
        self.nfe_step, self.cfg_strength, self.sway = nfe_step, cfg_strength, sway_sampling_coef
        self.max_mel_len = max_mel_len  # the engine's max sequence length (model.py:131)

    @staticmethod
    def _text(x) -> str:
        if isinstance(x, np.ndarray):
            x = x.reshape(-1)[0]
        return x.decode("utf-8") if isinstance(x, (bytes, bytearray)) else str(x)

    @torch.no_grad()
    def execute(self, requests: Sequence[dict]) -> List[np.ndarray]:
        if not requests:
            return []
        mels, rms_list, texts, est = [], [], [], []
        for r in requests:
            wav = torch.as_tensor(np.asarray(r["reference_wav"], dtype=np.float32)).reshape(1, -1)
            n = int(np.asarray(r["reference_wav_len"]).reshape(-1)[0]) if r.get("reference_wav_len") is not None else wav.shape[1]
            audio, rms = prepare_ref_audio((wav[:, :n], self.sr), self.target_rms, self.device)  # :205-212
            mel = self.model.mel_spec(audio).permute(0, 2, 1).to(torch.float32)  # [1, T, 100]
            ref_text, tgt_text = self._text(r["reference_text"]), self._text(r["target_text"])
            mels.append(mel[0])
            rms_list.append(rms)
            texts.append(ref_text + tgt_text)
            est.append(min(int(mel.shape[1] * (1 + len(tgt_text.encode("utf-8")) / len(ref_text.encode("utf-8")))),
                           self.max_mel_len))  # :222-229
        tokens = self.tokenize(texts)
        ragged_ok = bool(getattr(self.model.transformer, "attn_mask_enabled", False))
        groups = [list(range(len(requests)))] if ragged_ok or len(requests) == 1 else [[i] for i in range(len(requests))]
        out: List[Optional[np.ndarray]] = [None] * len(requests)
        for g in groups:
            b = len(g)
            ref_lens = torch.tensor([mels[i].shape[0] for i in g], dtype=torch.long)
            cond = torch.zeros((b, int(ref_lens.max()), mels[g[0]].shape[1]), device=self.device)
            for k, i in enumerate(g):
                cond[k, : mels[i].shape[0]] = mels[i]
            dur = torch.tensor([est[i] for i in g], dtype=torch.long)
            vocab = getattr(self.model, "vocab_char_map", None)
            text = list_str_to_idx([tokens[i] for i in g], vocab).to(self.device)
            mel_out, _ = self.model.sample(cond=cond, text=text, duration=dur.to(self.device) if b > 1 else int(dur[0]),
                                           lens=ref_lens.to(self.device), steps=self.nfe_step,
                                           cfg_strength=self.cfg_strength, sway_sampling_coef=self.sway,
                                           return_trajectory=False)
            # effective row length after cfm.py:135-137 (text / prompt longer than the estimate)
            eff = [max(est[i], max(int((text[k] != -1).sum()), int(ref_lens[k])) + 1) for k, i in enumerate(g)]
            glen = torch.tensor([min(eff[k], est[i]) - int(ref_lens[k]) for k, i in enumerate(g)], dtype=torch.int32)
            gen = torch.zeros((b, mel_out.shape[2], int(glen.max())), device=self.device)
            for k in range(b):  # denoised[i, ref_mel_len:estimated_mel_len] (:254-256)
                gen[k, :, : int(glen[k])] = mel_out[k, int(ref_lens[k]): int(ref_lens[k]) + int(glen[k])].T
            wav = self.vocoder.decode(gen, glen.to(self.device)) if b > 1 else self.vocoder.decode(gen)
            wav_h = wav.cpu().numpy()
            for k, i in enumerate(g):
                w = wav_h[k, : (int(glen[k]) - 1) * hop_length]
                if rms_list[i] < self.target_rms:  # :258-259
                    w = w * (rms_list[i] / self.target_rms)
                out[i] = w.astype(np.float32)
        return out  # type: ignore[return-value]
