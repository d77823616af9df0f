"""MelSpec with the reference's constructor (reference src/f5_tts/model/modules.py:107-146), computing on libf5b200's
hand-written STFT + mel-filterbank kernel (`f5b_mel`) instead of torchaudio / cuFFT."""
from __future__ import annotations

import torch
from torch import nn

from . import _lib


class MelSpec(nn.Module):
    def __init__(self, n_fft=1024, hop_length=256, win_length=1024, n_mel_channels=100, target_sample_rate=24_000,
                 mel_spec_type="vocos"):
        super().__init__()
        assert mel_spec_type in ["vocos", "bigvgan"], "We only support two extract mel backend: vocos or bigvgan"
        if mel_spec_type != "vocos":
            raise ValueError("the bigvgan mel variant is out of scope (north star names Vocos; SURVEY.md 2 #23)")
        if (n_fft, hop_length, win_length, n_mel_channels, target_sample_rate) != (1024, 256, 1024, 100, 24_000):
            raise ValueError("f5b_mel is built for the shipped mel_spec config: 24 kHz, n_fft 1024, hop 256, 100 mels")
        self.n_fft = n_fft
        self.hop_length = hop_length
        self.win_length = win_length
        self.n_mel_channels = n_mel_channels
        self.target_sample_rate = target_sample_rate
        self.device_index = None  # CUDA device used when the waveform arrives in host memory
        self.register_buffer("dummy", torch.tensor(0), persistent=False)

    def forward(self, wav: torch.Tensor) -> torch.Tensor:
        """wav [b, nw] (or [b, 1, nw]) -> log-mel [b, 100, nw // 256 + 1]."""
        if wav.ndim == 3:
            wav = wav.squeeze(1)
        assert wav.ndim == 2
        w = wav.to(torch.float32).contiguous()
        b, nw = w.shape
        frames = nw // self.hop_length + 1
        out = torch.empty((b, self.n_mel_channels, frames), device=w.device, dtype=torch.float32)
        if w.is_cuda:
            dev = _lib.device_index(w.device)
        elif self.dummy.is_cuda:
            dev = _lib.device_index(self.dummy.device)
        else:
            dev = self.device_index if self.device_index is not None else _lib.device_index(torch.device("cuda"))
        _lib.check(_lib.lib().f5b_mel(w.data_ptr(), b, nw, out.data_ptr(), dev, _lib.mem_kind(w),
                                      _lib.current_stream_ptr(w.device)))
        return out
