"""Host-side tensor helpers of the sampling path.  These stay PyTorch on the host/device the caller uses and must
match the reference exactly: reference src/f5_tts/model/utils.py:29-58,92-106,538-551."""
from __future__ import annotations

import random

import torch
from torch.nn.utils.rnn import pad_sequence

EPSS_GRIDS = {  # model/utils.py:540-547, in units of 1/32
    5: (0, 2, 4, 8, 16, 32),
    6: (0, 2, 4, 6, 8, 16, 32),
    7: (0, 2, 4, 6, 8, 16, 24, 32),
    10: (0, 2, 4, 6, 8, 12, 16, 20, 24, 28, 32),
    12: (0, 2, 4, 6, 8, 10, 12, 14, 16, 20, 24, 28, 32),
    16: (0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 14, 16, 20, 24, 28, 32),
}


def exists(v):
    return v is not None


def default(v, d):
    return v if v is not None else d


def seed_everything(seed=0):
    """model/utils.py:19-26."""
    import os

    random.seed(seed)
    os.environ["PYTHONHASHSEED"] = str(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(seed)
        torch.cuda.manual_seed_all(seed)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False


def lens_to_mask(t: torch.Tensor, length: int | None = None) -> torch.Tensor:
    """bool [b, n]: position < length of that row (model/utils.py:53-58)."""
    if length is None:
        length = t.amax()
    return torch.arange(length, device=t.device)[None, :] < t[:, None]


def list_str_to_tensor(text: list[str], padding_value=-1) -> torch.Tensor:
    """UTF-8 byte tokenizer (model/utils.py:92-95)."""
    rows = [torch.tensor([*bytes(t, "UTF-8")]) for t in text]
    return pad_sequence(rows, padding_value=padding_value, batch_first=True)


def list_str_to_idx(text, vocab_char_map: dict, padding_value=-1) -> torch.Tensor:
    """Vocabulary lookup, unknown symbol -> 0, right-padded with -1 (model/utils.py:99-106)."""
    rows = [torch.tensor([vocab_char_map.get(c, 0) for c in t]) for t in text]
    return pad_sequence(rows, padding_value=padding_value, batch_first=True)


def get_epss_timesteps(n: int, device, dtype) -> torch.Tensor:
    """Empirically pruned step grid for low NFE, uniform otherwise (model/utils.py:538-551)."""
    grid = EPSS_GRIDS.get(n)
    if grid is None:
        return torch.linspace(0, 1, n + 1, device=device, dtype=dtype)
    return (1 / 32) * torch.tensor(grid, device=device, dtype=dtype)


def sway_time_grid(steps: int, sway_sampling_coef, device, dtype, use_epss=True, t_start=0.0) -> torch.Tensor:
    """cfm.py:211-216: EPSS (or linspace) grid followed by the sway warp t + s*(cos(pi t/2) - 1 + t)."""
    if t_start == 0 and use_epss:
        t = get_epss_timesteps(steps, device=device, dtype=dtype)
    else:
        t = torch.linspace(t_start, 1, steps + 1, device=device, dtype=dtype)
    if sway_sampling_coef is not None:
        t = t + sway_sampling_coef * (torch.cos(torch.pi / 2 * t) - 1 + t)
    return t
