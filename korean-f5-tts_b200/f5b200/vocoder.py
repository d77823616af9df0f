"""Vocos (mel -> waveform) on libf5b200 and the reference's `load_vocoder` entry point.

Replaces the third-party `vocos.Vocos` object the reference builds in src/f5_tts/infer/utils_infer.py:114-137 and calls
as `vocoder.decode(mel)` at :703 and model/cfm.py:225-227.  Parameter names follow the vocos 0.1.0 checkpoint layout
(`charactr/vocos-mel-24khz`: backbone.embed / norm / convnext.{i}.* / final_layer_norm, head.out) so a released
`pytorch_model.bin` loads strictly; `feature_extractor.*` and `head.istft.window` buffers are accepted and ignored.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
from torch import nn

from . import _lib

VOCOS_CFG = dict(n_mels=100, dim=512, inter=1536, layers=8, n_fft=1024, hop=256)


class Vocos(nn.Module):
    def __init__(self, precision: str = "fp32"):
        super().__init__()
        c = VOCOS_CFG
        d, it = c["dim"], c["inter"]
        bb = nn.Module()
        bb.embed = nn.Conv1d(c["n_mels"], d, kernel_size=7, padding=3)
        bb.norm = nn.LayerNorm(d, eps=1e-6)
        blocks = []
        for _ in range(c["layers"]):
            m = nn.Module()
            m.dwconv = nn.Conv1d(d, d, kernel_size=7, padding=3, groups=d)
            m.norm = nn.LayerNorm(d, eps=1e-6)
            m.pwconv1 = nn.Linear(d, it)
            m.pwconv2 = nn.Linear(it, d)
            m.gamma = nn.Parameter(torch.full((d,), 1.0 / c["layers"]))
            blocks.append(m)
        bb.convnext = nn.ModuleList(blocks)
        bb.final_layer_norm = nn.LayerNorm(d, eps=1e-6)
        self.backbone = bb
        self.head = nn.Module()
        self.head.out = nn.Linear(d, c["n_fft"] + 2)
        self.precision = precision
        self._handle = None
        self._sig = None

    def load_state_dict(self, state_dict, strict: bool = True, **kw):
        keep = {k: v for k, v in state_dict.items()
                if not (k.startswith("feature_extractor.") or k == "head.istft.window")}
        return super().load_state_dict(keep, strict=strict, **kw)

    def _signature(self):
        ps = getattr(self, "_ps", None)  # cached Parameter list, dropped by _apply (see backbones._signature)
        if ps is None:
            ps = self._ps = list(self.parameters())
        return (str(ps[0].device), self.precision, tuple(p._version for p in ps), tuple(p.data_ptr() for p in ps))

    def _apply(self, fn, *args, **kwargs):
        self._ps = None
        return super()._apply(fn, *args, **kwargs)

    def _drop(self):
        if getattr(self, "_handle", None):
            _lib.lib().f5b_vocos_destroy(self._handle)
        self._handle = None

    def __del__(self):
        try:
            self._drop()
        except Exception:
            pass

    def handle(self):
        sig = self._signature()
        if self._handle is not None and sig == self._sig:
            return self._handle
        self._drop()
        device = next(self.parameters()).device
        arr, keep = _lib.make_tensor_array(self.state_dict())
        h = C.c_void_p()
        mode = _lib.MODE_BF16 if self.precision == "bf16" else _lib.MODE_FP32
        _lib.check(_lib.lib().f5b_vocos_create(arr, len(arr), mode, _lib.device_index(device), C.byref(h)))
        del keep
        self._handle, self._sig = h, sig
        return h

    @torch.no_grad()
    def decode(self, mel: torch.Tensor, lens: torch.Tensor | None = None) -> torch.Tensor:
        """mel [b, 100, T] -> waveform [b, (T - 1) * 256].  `lens` [b] (extra): row i holds lens[i] <= T frames -- it is
        decoded exactly as if it were alone (the convolutions' zero padding sits at ITS end); row i of the result is
        valid up to (lens[i] - 1) * 256 samples and zero after.  On CUDA tensors the call returns once enqueued."""
        h = self.handle()
        m = mel.to(torch.float32).contiguous()
        b, c, t = m.shape
        assert c == VOCOS_CFG["n_mels"]
        wav = torch.empty((b, (t - 1) * VOCOS_CFG["hop"]), device=m.device, dtype=torch.float32)
        ln = None if lens is None else lens.to(device=m.device, dtype=torch.int32).contiguous()
        _lib.check(_lib.lib().f5b_vocos_decode_varlen(h, m.data_ptr(), _lib.ptr(ln), b, t, wav.data_ptr(),
                                                      _lib.mem_kind(m), _lib.current_stream_ptr(m.device)))
        return wav

    forward = decode

    def reserve(self, batch: int, frames: int):
        """Grow the workspace up front for a decode of `batch` x `frames` (monotone capacity: f5b_vocos_reserve)."""
        _lib.check(_lib.lib().f5b_vocos_reserve(self.handle(), int(batch), int(frames)))


def load_vocoder(vocoder_name="vocos", is_local=False, local_path="", device="cuda", hf_cache_dir=None, *,
                 precision: str | None = None, allow_uninitialized: bool = False):
    """Same signature as the reference (infer/utils_infer.py:113).  The reference downloads `charactr/vocos-mel-24khz`
    from the hub when `is_local` is False; there is no network here, so weights must come from `local_path`
    (`pytorch_model.bin`, or a hub cache laid out under `hf_cache_dir`).  Returning a randomly initialised vocoder would
    make `F5TTS.infer` emit noise without any error: that needs the explicit `allow_uninitialized=True`
    (benchmarks and tests that `load_state_dict` synthetic weights afterwards).
    `precision` None follows the backbone's rule: 16-bit tensor-core kernels on CUDA (the reference runs fp16 there,
    utils_infer.py:243-251), fp32 kernels when asked for."""
    if vocoder_name != "vocos":
        raise ValueError("only the Vocos vocoder is built (north star); BigVGAN is out of scope")
    if precision is None:
        precision = "bf16" if "cuda" in str(device) else "fp32"
    voc = Vocos(precision=precision)
    model_path = None
    if is_local and local_path:
        model_path = os.path.join(local_path, "pytorch_model.bin")
    elif hf_cache_dir:  # a pre-populated hub cache: <cache>/models--charactr--vocos-mel-24khz/snapshots/*/pytorch_model.bin
        import glob

        hits = sorted(glob.glob(os.path.join(hf_cache_dir, "models--charactr--vocos-mel-24khz", "snapshots", "*",
                                             "pytorch_model.bin")))
        model_path = hits[-1] if hits else None
    if model_path is not None:
        state = torch.load(model_path, map_location="cpu", weights_only=True)
        voc.load_state_dict(state)
    elif not allow_uninitialized:
        raise FileNotFoundError("load_vocoder: no Vocos weights (the reference's hub download needs a network): pass "
                                "is_local=True with local_path=<dir holding pytorch_model.bin>, or "
                                "allow_uninitialized=True and load a state dict yourself")
    return voc.eval().to(device)
