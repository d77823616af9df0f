"""Stand-ins for third-party modules the reference imports but this image lacks.

TEST INFRASTRUCTURE (see oracle/__init__.py).  These are the only pieces of
restated third-party arithmetic used when the *unmodified* reference files are
executed in this container:

* ``torchdiffeq.odeint(method="euler")``  (pyproject.toml:39, unpinned; call site
  src/f5_tts/model/cfm.py:20,218): fixed-grid Euler, returns the stacked trajectory.
* ``x_transformers.x_transformers.RotaryEmbedding / apply_rotary_pos_emb`` and
  ``x_transformers.RMSNorm`` (pyproject.toml:46, ">=1.31.14"; call sites dit.py:17,184,311,
  unett.py:18-19, modules.py:21,493-497).  Second witnesses inside the reference for the
  RoPE convention: runtime/triton_trtllm/.../f5_tts_trtllm.py:229-237 and
  patch/f5tts/modules.py:210-276.
* ``librosa.filters.mel`` (only used by the out-of-scope bigvgan mel), ``rjieba``, ``pypinyin``
  (module-level imports of model/utils.py:10-12; never executed on this path).

parity unpinned: the reference holds no tests for these third-party pieces.
"""
from __future__ import annotations

import math
import sys
import types

import torch
from torch import nn


def _euler_odeint(fn, y0, t, method="euler", **_):
    assert method == "euler", "only the fixed-grid Euler solver is restated"
    ys = [y0]
    y = y0
    for i in range(t.shape[0] - 1):
        dt = t[i + 1] - t[i]
        y = y + dt * fn(t[i], y)
        ys.append(y)
    return torch.stack(ys, dim=0)


class RotaryEmbedding(nn.Module):
    def __init__(self, dim, base=10000.0):
        super().__init__()
        inv_freq = 1.0 / (base ** (torch.arange(0, dim, 2).float() / dim))
        self.register_buffer("inv_freq", inv_freq)  # persistent, like x_transformers': checkpoints carry the key

    def forward_from_seq_len(self, seq_len):
        pos = torch.arange(seq_len, device=self.inv_freq.device).type_as(self.inv_freq)
        ang = pos[None, :, None] * self.inv_freq[None, None, :]  # [1, n, dim/2]
        ang = torch.stack((ang, ang), dim=-1).flatten(-2)  # pairwise duplicated -> [1, n, dim]
        return ang, 1.0


def _rotate_pairs(x):
    x = x.unflatten(-1, (-1, 2))
    a, b = x.unbind(dim=-1)
    return torch.stack((-b, a), dim=-1).flatten(-2)


def apply_rotary_pos_emb(t, freqs, scale=1):
    rot_dim, seq_len, orig_dtype = freqs.shape[-1], t.shape[-2], t.dtype
    freqs = freqs[:, -seq_len:, :]
    if t.ndim == 4 and freqs.ndim == 3:
        freqs = freqs[:, None]
    head, tail = t[..., :rot_dim], t[..., rot_dim:]
    head = (head * freqs.cos() * scale) + (_rotate_pairs(head) * freqs.sin() * scale)
    return torch.cat((head, tail), dim=-1).type(orig_dtype)


class XRMSNorm(nn.Module):
    """x_transformers.RMSNorm: F.normalize(x, dim=-1) * sqrt(dim) * g."""

    def __init__(self, dim):
        super().__init__()
        self.scale = dim**0.5
        self.g = nn.Parameter(torch.ones(dim))

    def forward(self, x):
        return torch.nn.functional.normalize(x, dim=-1) * self.scale * self.g


def install():
    """Register the stub modules in sys.modules (idempotent)."""
    if "torchdiffeq" not in sys.modules:
        m = types.ModuleType("torchdiffeq")
        m.odeint = _euler_odeint
        sys.modules["torchdiffeq"] = m
    if "x_transformers" not in sys.modules:
        xt = types.ModuleType("x_transformers")
        xt.__path__ = []
        xt.RMSNorm = XRMSNorm
        xx = types.ModuleType("x_transformers.x_transformers")
        xx.RotaryEmbedding = RotaryEmbedding
        xx.apply_rotary_pos_emb = apply_rotary_pos_emb
        xt.x_transformers = xx
        sys.modules["x_transformers"] = xt
        sys.modules["x_transformers.x_transformers"] = xx
    if "librosa" not in sys.modules:
        lb = types.ModuleType("librosa")
        lb.__path__ = []
        lf = types.ModuleType("librosa.filters")

        def _mel(*a, **k):
            raise NotImplementedError("bigvgan mel is out of scope (SURVEY.md §2 #23)")

        lf.mel = _mel
        lb.filters = lf
        sys.modules["librosa"] = lb
        sys.modules["librosa.filters"] = lf
    if "rjieba" not in sys.modules:
        rj = types.ModuleType("rjieba")
        rj.cut = lambda s: list(s)
        sys.modules["rjieba"] = rj
    if "pypinyin" not in sys.modules:
        pp = types.ModuleType("pypinyin")

        class Style:  # noqa: D401
            TONE3 = 0

        pp.Style = Style
        pp.lazy_pinyin = lambda s, **k: list(s)
        sys.modules["pypinyin"] = pp
