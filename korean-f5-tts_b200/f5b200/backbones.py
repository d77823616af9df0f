"""DiT and UNetT with the reference's constructor signatures, parameter names and shapes, executing on libf5b200.

The module tree below exists only to hold parameters (so `load_state_dict` of reference checkpoints is strict-
compatible: reference src/f5_tts/infer/utils_infer.py:275-281) and to give them the reference's initialisation
(dit.py:214-224 zero-inits the AdaLN linears and proj_out; modules.py:234-235 zero-inits GRN).  `forward` hands raw
pointers to the C ABI (`f5b_forward`), which replaces backbones/dit.py:278-329 and backbones/unett.py:217-280.
There is no PyTorch execution path.
"""
from __future__ import annotations

import ctypes as C

import torch
from torch import nn

from . import _lib


def _convnext_v2_block(dim: int, inter: int) -> nn.Module:
    m = nn.Module()
    m.dwconv = nn.Conv1d(dim, dim, kernel_size=7, padding=3, groups=dim)
    m.norm = nn.LayerNorm(dim, eps=1e-6)
    m.pwconv1 = nn.Linear(dim, inter)
    m.grn = nn.Module()
    m.grn.gamma = nn.Parameter(torch.zeros(1, 1, inter))
    m.grn.beta = nn.Parameter(torch.zeros(1, 1, inter))
    m.pwconv2 = nn.Linear(inter, dim)
    return m


def _text_embedding(text_num_embeds: int, text_dim: int, conv_layers: int) -> nn.Module:
    m = nn.Module()
    m.text_embed = nn.Embedding(text_num_embeds + 1, text_dim)
    if conv_layers > 0:
        m.text_blocks = nn.Sequential(*[_convnext_v2_block(text_dim, text_dim * 2) for _ in range(conv_layers)])
    return m


def _input_embedding(mel_dim: int, text_dim: int, dim: int) -> nn.Module:
    m = nn.Module()
    m.proj = nn.Linear(mel_dim * 2 + text_dim, dim)
    m.conv_pos_embed = nn.Module()
    m.conv_pos_embed.conv1d = nn.Sequential(
        nn.Conv1d(dim, dim, 31, groups=16, padding=15), nn.Mish(), nn.Conv1d(dim, dim, 31, groups=16, padding=15),
        nn.Mish())
    return m


def _time_embedding(dim: int) -> nn.Module:
    m = nn.Module()
    m.time_mlp = nn.Sequential(nn.Linear(256, dim), nn.SiLU(), nn.Linear(dim, dim))
    return m


def _attention(dim: int, heads: int, dim_head: int, dropout: float) -> nn.Module:
    inner = heads * dim_head
    m = nn.Module()
    m.to_q = nn.Linear(dim, inner)
    m.to_k = nn.Linear(dim, inner)
    m.to_v = nn.Linear(dim, inner)
    m.to_out = nn.ModuleList([nn.Linear(inner, dim), nn.Dropout(dropout)])
    return m


def _feed_forward(dim: int, mult: int, dropout: float) -> nn.Module:
    inner = int(dim * mult)
    m = nn.Module()
    m.ff = nn.Sequential(nn.Sequential(nn.Linear(dim, inner), nn.GELU(approximate="tanh")), nn.Dropout(dropout),
                         nn.Linear(inner, dim))
    return m


def _x_rmsnorm(dim: int) -> nn.Module:
    m = nn.Module()
    m.g = nn.Parameter(torch.ones(dim))
    return m


class _RotaryEmbedding(nn.Module):
    """Holds `inv_freq` under the name the reference's state dicts carry it: x_transformers' RotaryEmbedding registers
    it as a PERSISTENT buffer (dit.py:184, unett.py:140), so trainer-written checkpoints contain
    `transformer.rotary_embed.inv_freq` and a strict load needs a home for it.  The library computes its own RoPE table
    (the same fp32 expression); a checkpoint without the key (written by code that keeps the buffer non-persistent) loads
    as well."""

    def __init__(self, dim_head: int, base: float = 10000.0):
        super().__init__()
        self.register_buffer("inv_freq", 1.0 / (base ** (torch.arange(0, dim_head, 2).float() / dim_head)))

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        if prefix + "inv_freq" not in state_dict:
            return
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)


class _PlanBackbone(nn.Module):
    """Shared plan lifecycle: parameters are fp32 masters; `precision` selects the kernel set."""

    _backbone_id = _lib.BACKBONE_DIT

    def _init_common(self, precision: str):
        self._plan = None
        self._plan_sig = None
        self.precision = precision

    # --- precision handling: .half()/.bfloat16()/.to(dtype) choose the bf16 tensor-core kernels, .float() the fp32 set
    def to(self, *args, **kwargs):
        device, dtype, non_blocking, _ = torch._C._nn._parse_to(*args, **kwargs)
        if dtype is not None:
            if dtype in (torch.float16, torch.bfloat16):
                self.precision = "bf16"
            elif dtype == torch.float32:
                self.precision = "fp32"
            else:
                raise ValueError(f"unsupported dtype {dtype}")
        if device is not None:
            super().to(device=device, non_blocking=non_blocking)
        return self

    def half(self):
        self.precision = "bf16"
        return self

    def bfloat16(self):
        self.precision = "bf16"
        return self

    def float(self):
        self.precision = "fp32"
        return self

    def _arch(self) -> _lib.Arch:
        raise NotImplementedError

    def _signature(self):
        # Walking the module tree costs ~1 ms of Python per call (363 parameters); the list of Parameter objects is kept
        # and dropped whenever nn.Module._apply runs (.to / .cuda / .float ...).  In-place updates (load_state_dict,
        # copy_, optimizer steps) keep the objects and bump `_version`, which the signature still sees.
        ps = getattr(self, "_ps", None)
        if ps is None:
            ps = self._ps = list(self.parameters())
        return (str(ps[0].device), self.precision, tuple(p._version for p in ps), tuple(p.data_ptr() for p in ps))

    def _apply(self, fn, *args, **kwargs):
        self._ps = None
        return super()._apply(fn, *args, **kwargs)

    def invalidate_plan(self):
        """Call after replacing a Parameter object of a sub-module (in-place updates are detected automatically)."""
        self._ps = None
        self._drop_plan()

    def _drop_plan(self):
        if getattr(self, "_plan", None):
            _lib.lib().f5b_plan_destroy(self._plan)
        self._plan = None
        self._plan_sig = None

    def __del__(self):
        try:
            self._drop_plan()
        except Exception:
            pass

    def plan(self):
        """Create (or reuse) the device plan for the current weights / device / precision."""
        sig = self._signature()
        if self._plan is not None and sig == self._plan_sig:
            return self._plan
        self._drop_plan()
        device = next(self.parameters()).device
        dev_index = _lib.device_index(device)
        arr, keep = _lib.make_tensor_array(self.state_dict())
        arch = self._arch()
        handle = C.c_void_p()
        mode = _lib.MODE_BF16 if self.precision == "bf16" else _lib.MODE_FP32
        _lib.check(_lib.lib().f5b_plan_create(C.byref(arch), arr, len(arr), mode, dev_index, C.byref(handle)))
        del keep
        self._plan = handle
        self._plan_sig = sig
        return handle

    def reserve(self, shapes, steps: int = 32):
        """Grow the workspace up front for every (batch, frames, text tokens) shape of a job, so that no call inside the
        job re-allocates (capacities are per quantity and monotone: f5b_plan_reserve)."""
        plan = self.plan()
        for b, n, nt in shapes:
            _lib.check(_lib.lib().f5b_plan_reserve(plan, int(b), int(n), int(max(nt, 1)), int(steps)))

    def refresh_plan(self):
        """After parameters were updated in place behind torch's back (TrainStep's fused AdamW writes the flat buffer
        the parameters are views of): re-pack the weights into the existing plan (no allocation, graphs stay valid)."""
        if getattr(self, "_plan", None) is None:
            return self.plan()
        arr, keep = _lib.make_tensor_array(self.state_dict())
        dev = next(self.parameters()).device
        _lib.check(_lib.lib().f5b_plan_refresh(self._plan, arr, len(arr), _lib.current_stream_ptr(dev)))
        del keep
        self._plan_sig = self._signature()
        return self._plan

    def synchronize(self):
        """Wait for the plan's last call and raise its deferred error.  Calls on CUDA tensors return once their work is
        enqueued (like any PyTorch CUDA operator); a token id outside the embedding table -- `nn.Embedding` would raise
        IndexError (dit.py:42,94) -- then surfaces here, or in the next call on this module."""
        if self._plan is not None:
            _lib.check(_lib.lib().f5b_plan_sync(self._plan))

    def clear_cache(self):
        if self._plan is not None:
            _lib.lib().f5b_clear_cache(self._plan)

    def forward(self, x, cond, text, time, mask=None, drop_audio_cond=False, drop_text=False, cfg_infer=False,
                cache=False):
        plan = self.plan()
        dev = x.device
        b, n, mel = x.shape
        x32 = x.to(torch.float32).contiguous()
        c32 = cond.to(torch.float32).contiguous()
        tx = text.to(device=dev, dtype=torch.long).contiguous()
        time = torch.as_tensor(time, dtype=torch.float32)
        if time.ndim == 0:
            time = time.repeat(b)
        t_host = time.detach().to("cpu", torch.float32).contiguous()
        dur = None
        if mask is not None:
            dur = mask.sum(dim=-1).to(device=dev, dtype=torch.int32).contiguous()
        out = torch.empty((2 * b if cfg_infer else b, n, mel), device=dev, dtype=torch.float32)
        _lib.check(_lib.lib().f5b_forward(
            plan, x32.data_ptr(), c32.data_ptr(), tx.data_ptr(), t_host.data_ptr(), _lib.ptr(dur), b, n, tx.shape[1],
            int(drop_audio_cond), int(drop_text), int(cfg_infer), int(cache), out.data_ptr(), _lib.mem_kind(x32),
            _lib.current_stream_ptr(dev)))
        return out.to(x.dtype)


class DiT(_PlanBackbone):
    """Constructor mirrors reference backbones/dit.py:147-168."""

    _backbone_id = _lib.BACKBONE_DIT

    def __init__(self, *, dim, depth=8, heads=8, dim_head=64, dropout=0.1, ff_mult=4, mel_dim=100,
                 text_num_embeds=256, text_dim=None, text_mask_padding=True, text_embedding_average_upsampling=False,
                 qk_norm=None, conv_layers=0, pe_attn_head=None, attn_backend="torch", attn_mask_enabled=False,
                 long_skip_connection=False, checkpoint_activations=False, precision="bf16"):
        super().__init__()
        if text_embedding_average_upsampling or qk_norm is not None or long_skip_connection:
            raise ValueError("text_embedding_average_upsampling / qk_norm / long_skip_connection are not used by any "
                             "shipped config and are not built (DESIGN.md, out of scope)")
        if text_dim is None:
            text_dim = mel_dim
        self._init_common(precision)
        self.dim, self.depth, self.heads, self.dim_head, self.ff_mult = dim, depth, heads, dim_head, ff_mult
        self.mel_dim, self.text_num_embeds, self.text_dim = mel_dim, text_num_embeds, text_dim
        self.text_mask_padding, self.conv_layers = text_mask_padding, conv_layers
        self.pe_attn_head, self.attn_mask_enabled = pe_attn_head, attn_mask_enabled
        self.time_embed = _time_embedding(dim)
        self.text_embed = _text_embedding(text_num_embeds, text_dim, conv_layers)
        self.input_embed = _input_embedding(mel_dim, text_dim, dim)
        blocks = []
        for _ in range(depth):
            blk = nn.Module()
            blk.attn_norm = nn.Module()
            blk.attn_norm.linear = nn.Linear(dim, dim * 6)
            blk.attn = _attention(dim, heads, dim_head, dropout)
            blk.ff = _feed_forward(dim, ff_mult, dropout)
            nn.init.constant_(blk.attn_norm.linear.weight, 0)  # dit.py:216-218
            nn.init.constant_(blk.attn_norm.linear.bias, 0)
            blocks.append(blk)
        self.rotary_embed = _RotaryEmbedding(dim_head)
        self.transformer_blocks = nn.ModuleList(blocks)
        self.norm_out = nn.Module()
        self.norm_out.linear = nn.Linear(dim, dim * 2)
        self.proj_out = nn.Linear(dim, mel_dim)
        for p in (self.norm_out.linear.weight, self.norm_out.linear.bias, self.proj_out.weight, self.proj_out.bias):
            nn.init.constant_(p, 0)  # dit.py:221-224

    def _arch(self):
        return _lib.Arch(self._backbone_id, self.dim, self.depth, self.heads, self.dim_head, self.ff_mult, self.mel_dim,
                         self.text_num_embeds, self.text_dim, self.conv_layers, int(self.text_mask_padding),
                         self.pe_attn_head if self.pe_attn_head else 0, int(self.attn_mask_enabled))


class UNetT(_PlanBackbone):
    """Constructor mirrors reference backbones/unett.py:108-127."""

    _backbone_id = _lib.BACKBONE_UNETT

    def __init__(self, *, dim, depth=8, heads=8, dim_head=64, dropout=0.1, ff_mult=4, mel_dim=100,
                 text_num_embeds=256, text_dim=None, text_mask_padding=True, qk_norm=None, conv_layers=0,
                 pe_attn_head=None, attn_backend="torch", attn_mask_enabled=False, skip_connect_type="concat",
                 precision="bf16"):
        super().__init__()
        assert skip_connect_type in ["add", "concat", "none"]  # unett.py:129
        if skip_connect_type != "concat" or qk_norm is not None:
            raise ValueError("only skip_connect_type='concat' without qk_norm is built (E2TTS_Base.yaml)")
        if text_dim is None:
            text_dim = mel_dim
        self._init_common(precision)
        self.dim, self.depth, self.heads, self.dim_head, self.ff_mult = dim, depth, heads, dim_head, ff_mult
        self.mel_dim, self.text_num_embeds, self.text_dim = mel_dim, text_num_embeds, text_dim
        self.text_mask_padding, self.conv_layers = text_mask_padding, conv_layers
        self.pe_attn_head, self.attn_mask_enabled = pe_attn_head, attn_mask_enabled
        self.time_embed = _time_embedding(dim)
        self.text_embed = _text_embedding(text_num_embeds, text_dim, conv_layers)
        self.input_embed = _input_embedding(mel_dim, text_dim, dim)
        layers = []
        for idx in range(depth):
            later = (idx + 1) > (depth // 2)
            skip = nn.Linear(dim * 2, dim, bias=False) if later else None
            layers.append(nn.ModuleList([skip, _x_rmsnorm(dim), _attention(dim, heads, dim_head, dropout),
                                         _x_rmsnorm(dim), _feed_forward(dim, ff_mult, dropout)]))
        self.rotary_embed = _RotaryEmbedding(dim_head)
        self.layers = nn.ModuleList(layers)
        self.norm_out = _x_rmsnorm(dim)
        self.proj_out = nn.Linear(dim, mel_dim)

    def _arch(self):
        return _lib.Arch(self._backbone_id, self.dim, self.depth, self.heads, self.dim_head, self.ff_mult, self.mel_dim,
                         self.text_num_embeds, self.text_dim, self.conv_layers, int(self.text_mask_padding),
                         self.pe_attn_head if self.pe_attn_head else 0, int(self.attn_mask_enabled))
