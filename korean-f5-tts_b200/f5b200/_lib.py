"""ctypes binding of libf5b200.so (include/f5b200.h).  There is no Python/CPU fallback: if the shared library is
missing or no sm_100 device is present, every compute call raises."""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterable, Mapping

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libf5b200.so")

MEM_HOST, MEM_DEVICE = 0, 1
MODE_FP32, MODE_BF16 = 0, 1
BACKBONE_DIT, BACKBONE_UNETT = 0, 1
ACT = {"none": 0, "gelu_tanh": 1, "gelu": 2, "mish": 3, "silu": 4}

ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_MISSING, ERR_INDEX = -1, -2, -3, -4, -5


class F5BError(RuntimeError):
    pass


class Arch(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "backbone", "dim", "depth", "heads", "dim_head", "ff_mult", "mel_dim", "text_num_embeds", "text_dim",
        "conv_layers", "text_mask_padding", "pe_attn_head", "attn_mask_enabled")]


class Tensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data", C.c_void_p), ("numel", C.c_int64), ("on_device", C.c_int32)]


class Grad(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data", C.c_void_p), ("numel", C.c_int64), ("ready_event", C.c_void_p)]


_SIGNATURES = {
    "f5b_last_error": (C.c_char_p, []),
    "f5b_version": (C.c_int, []),
    "f5b_device_check": (C.c_int, [C.c_int]),
    "f5b_plan_create": (C.c_int, [C.POINTER(Arch), C.POINTER(Tensor), C.c_int32, C.c_int32, C.c_int32,
                                  C.POINTER(C.c_void_p)]),
    "f5b_plan_destroy": (None, [C.c_void_p]),
    "f5b_sample": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                             C.c_int32, C.c_void_p, C.c_int32, C.c_float, C.c_void_p, C.c_void_p, C.c_int32,
                             C.c_void_p]),
    "f5b_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                              C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32,
                              C.c_void_p]),
    "f5b_clear_cache": (C.c_int, [C.c_void_p]),
    "f5b_plan_sync": (C.c_int, [C.c_void_p]),
    "f5b_mel": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "f5b_ref_audio": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                C.c_float, C.c_void_p, C.c_int64, C.POINTER(C.c_float), C.c_int32, C.c_int32,
                                C.c_void_p]),
    "f5b_vocos_create": (C.c_int, [C.POINTER(Tensor), C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "f5b_vocos_destroy": (None, [C.c_void_p]),
    "f5b_vocos_decode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]),
    "f5b_vocos_decode_varlen": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32,
                                          C.c_void_p]),
    "f5b_op_istft": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]),
    "f5b_op_gemm": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                              C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "f5b_op_attention": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                   C.c_int32, C.c_void_p]),
    "f5b_plan_reserve": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "f5b_vocos_reserve": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "f5b_plan_refresh": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "f5b_op_attention_bwd": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                       C.c_void_p, C.c_int32, C.c_void_p]),
    "f5b_cfm_loss": (C.c_int, [C.c_void_p] * 7 + [C.c_int32] * 5 + [C.c_void_p] * 3 + [C.c_int32, C.c_void_p]),
    "f5b_cfm_loss_tail_grads": (C.c_int, [C.c_void_p] * 7 + [C.c_int32] * 5 + [C.c_void_p, C.POINTER(Grad), C.c_int32,
                                                                        C.c_int32, C.c_void_p]),
    "f5b_op_sumsq": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "f5b_op_sumsq_static": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p]),
    "f5b_op_adamw": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float,
                               C.c_float, C.c_float, C.c_float, C.c_int32, C.c_float, C.c_float, C.c_int32, C.c_void_p]),
    "f5b_op_time": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_float), C.c_int32,
                              C.c_void_p]),
    "f5b_debug_attn_trace": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32]),
    "f5b_debug_ktrace": (C.c_int, [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32]),
    "f5b_launch_count": (C.c_int64, []),
    "f5b_last_call_launches": (C.c_int64, []),
}

EXPORTED = tuple(_SIGNATURES)
_lib = None


def lib() -> C.CDLL:
    """Load the library once.  Raises F5BError (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise F5BError(f"{LIB_PATH} not built: run `python korean-f5-tts_b200/build.py` (no CPU fallback exists)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(code: int) -> None:
    if code == 0:
        return
    msg = lib().f5b_last_error().decode("utf-8", "replace")
    if code == ERR_INVALID:
        raise ValueError(msg)
    if code == ERR_MISSING:
        raise KeyError(msg)
    if code == ERR_INDEX:
        raise IndexError(msg)
    raise F5BError(f"libf5b200 error {code}: {msg}")


def current_stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream if device.type == "cuda" else 0


def ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def mem_kind(t: torch.Tensor) -> int:
    return MEM_DEVICE if t.is_cuda else MEM_HOST


def make_tensor_array(state: Mapping[str, torch.Tensor], prefix_strip: Iterable[str] = ()):
    """state dict -> (ctypes array of f5b_tensor, keep-alive list).  Tensors are made fp32 + contiguous."""
    keep = []
    items = []
    for name, t in state.items():
        for p in prefix_strip:
            if name.startswith(p):
                name = name[len(p):]
        t = t.detach().to(torch.float32).contiguous()
        if t.numel() == 0:
            continue
        bname = name.encode()
        keep.append((bname, t))
        items.append(Tensor(bname, t.data_ptr(), t.numel(), 1 if t.is_cuda else 0))
    arr = (Tensor * len(items))(*items)
    return arr, keep


def device_index(device: torch.device) -> int:
    if device.type != "cuda":
        raise F5BError("libf5b200 computes on a CUDA sm_100 device only; move the module to cuda (no CPU fallback)")
    if device.index is not None:
        return device.index
    return torch.cuda.current_device() if torch.cuda.is_available() else 0
