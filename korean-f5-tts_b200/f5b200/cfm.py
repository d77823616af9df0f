"""CFM with the reference's constructor / `sample` signature (reference src/f5_tts/model/cfm.py:34-229).

Host side (PyTorch, must match the reference exactly): mel of a raw-wave prompt, lens / duration / mask
construction, token lookup, seeded noise, EPSS + sway time grid, the final `where(cond_mask, cond, out)`.
Device side (libf5b200, one call): text embedding cache, every CFG-doubled backbone evaluation, Euler updates.
`forward` (cfm.py:231-302) evaluates the flow-matching loss through one C-ABI call; it produces no gradients.
"""
from __future__ import annotations

from random import random
from typing import Callable

import torch
import torch.nn.functional as F
from torch import nn
from torch.nn.utils.rnn import pad_sequence

from . import _lib
from .modules import MelSpec
from .utils import default, exists, lens_to_mask, list_str_to_idx, list_str_to_tensor, sway_time_grid


class CFM(nn.Module):
    def __init__(self, transformer: nn.Module, sigma=0.0, odeint_kwargs: dict = dict(method="euler"),
                 audio_drop_prob=0.3, cond_drop_prob=0.2, num_channels=None, mel_spec_module: nn.Module | None = None,
                 mel_spec_kwargs: dict = dict(), frac_lengths_mask: tuple[float, float] = (0.7, 1.0),
                 vocab_char_map: dict | None = None):
        super().__init__()
        self.frac_lengths_mask = frac_lengths_mask
        self.mel_spec = default(mel_spec_module, MelSpec(**mel_spec_kwargs))
        self.num_channels = default(num_channels, self.mel_spec.n_mel_channels)
        self.audio_drop_prob = audio_drop_prob
        self.cond_drop_prob = cond_drop_prob
        self.transformer = transformer
        self.dim = transformer.dim
        self.sigma = sigma
        if odeint_kwargs.get("method", "euler") != "euler":
            raise ValueError("only the fixed-grid Euler solver is built (the reference's default, cfm.py:39-43)")
        self.odeint_kwargs = odeint_kwargs
        self.vocab_char_map = vocab_char_map

    @property
    def device(self):
        return next(self.parameters()).device

    def _prepare_loss(self, inp, text, lens, draws):
        """Host half of CFM.forward (cfm.py:231-282): the random draws in the reference's order and on its devices
        (`frac_lengths`, the span start, `x0`, `time` on `self.device`; the two drop decisions from `random()`), the
        span mask (model/utils.py:61-77), contiguous fp32 / int tensors for the C ABI."""
        if inp.ndim == 2:  # raw wave (cfm.py:240-243)
            inp = self.mel_spec(inp).permute(0, 2, 1)
            assert inp.shape[-1] == self.num_channels
        dev = self.device
        inp = inp.to(dev, torch.float32)
        batch, seq_len = inp.shape[:2]
        if isinstance(text, list):  # cfm.py:248-253
            text = (list_str_to_idx(text, self.vocab_char_map) if exists(self.vocab_char_map)
                    else list_str_to_tensor(text)).to(dev)
            assert text.shape[0] == batch
        if not exists(lens):
            lens = torch.full((batch,), seq_len, device=dev)
        lens = lens.to(dev)
        mask = lens_to_mask(lens, length=seq_len)
        if draws is None:
            frac_lengths = torch.zeros((batch,), device=dev).float().uniform_(*self.frac_lengths_mask)
            rand = torch.rand_like(frac_lengths)
            x0 = torch.randn_like(inp)
            time = torch.rand((batch,), dtype=inp.dtype, device=dev)
            drop_audio_cond = random() < self.audio_drop_prob
            if random() < self.cond_drop_prob:
                drop_audio_cond, drop_text = True, True
            else:
                drop_text = False
        else:
            frac_lengths, rand = draws["frac_lengths"].to(dev), draws["rand"].to(dev)
            x0, time = draws["x0"].to(dev, torch.float32), draws["time"].to(dev, torch.float32)
            drop_audio_cond, drop_text = bool(draws["drop_audio_cond"]), bool(draws["drop_text"])
        span_len = (frac_lengths * lens).long()
        start = ((lens - span_len) * rand).long().clamp(min=0)
        seq = torch.arange(seq_len, device=dev)
        span = (seq[None, :] >= start[:, None]) & (seq[None, :] < (start + span_len)[:, None]) & mask
        return dict(x1=inp.contiguous(), x0=x0.contiguous(), tx=text.to(device=dev, dtype=torch.long).contiguous(),
                    ln=lens.to(torch.int32).contiguous(), sp=span.to(torch.uint8).contiguous(),
                    t_host=time.detach().to("cpu", torch.float32).contiguous(), batch=batch, seq_len=seq_len,
                    drop_audio_cond=drop_audio_cond, drop_text=drop_text, dev=dev)

    def forward(self, inp, text, *, lens=None, noise_scheduler=None, draws: dict | None = None):
        """Flow-matching loss of one batch (cfm.py:231-302) -> (loss, cond, pred): the host draws (`_prepare_loss`), then
        phi, the infilling condition, the backbone evaluation and the masked MSE in one `f5b_cfm_loss` call.
        `draws` (keys frac_lengths, rand, x0, time, drop_audio_cond, drop_text) injects the random quantities (parity
        tests).  The result carries no autograd graph; gradients come from `tail_gradients` (the backward pass)."""
        q = self._prepare_loss(inp, text, lens, draws)
        x1, dev = q["x1"], q["dev"]
        loss = torch.empty((1,), device=dev, dtype=torch.float32)
        cond = torch.empty_like(x1)
        pred = torch.empty_like(x1)
        _lib.check(_lib.lib().f5b_cfm_loss(
            self.transformer.plan(), x1.data_ptr(), q["x0"].data_ptr(), q["tx"].data_ptr(), q["t_host"].data_ptr(),
            q["ln"].data_ptr(), q["sp"].data_ptr(), q["batch"], q["seq_len"], q["tx"].shape[1],
            int(q["drop_audio_cond"]), int(q["drop_text"]), loss.data_ptr(), cond.data_ptr(), pred.data_ptr(),
            _lib.mem_kind(x1), _lib.current_stream_ptr(dev)))
        return loss[0], cond, pred

    def tail_gradients(self, inp, text, *, lens=None, draws: dict | None = None, names=None, buffers: dict | None = None,
                       events: dict | None = None):
        """The backward pass (trainer.py:384 over CFM.forward): the loss and its gradients with respect to EVERY parameter
        of the DiT -- `{state-dict name: tensor}` with the reference's names (`transformer.` prefix included) -- plus
        `d_block_input` / `d_input_embed` (gradient of the residual stream entering the last block / block 0).  One
        `f5b_cfm_loss_tail_grads` call; CUDA-core arithmetic in fp32 mode, tcgen05 kernels in 16-bit mode; blocks recomputed
        from their saved inputs (activation checkpointing); dropout is not modelled (eval-mode function).  `buffers`
        ({name: preallocated fp32 tensor}) receives the gradients in place (the training step's flat gradient buffer);
        `events` ({name: cudaEvent_t handle}) are recorded by the library as each gradient lands (TrainStep's buckets)."""
        import ctypes as C

        tr = self.transformer
        q = self._prepare_loss(inp, text, lens, draws)
        x1, dev = q["x1"], q["dev"]
        params = dict(tr.named_parameters())
        if names is None:
            names = list(params)  # every parameter of the DiT
            names += ["d_block_input", "d_input_embed"]
        out, slots, keep = {}, [], []
        for n in names:
            shape = (q["batch"], q["seq_len"], tr.dim) if n in ("d_block_input", "d_input_embed") else \
                (q["batch"], q["seq_len"], self.num_channels) if n == "pred" else tuple(params[n].shape)
            g = buffers[n] if buffers is not None and n in buffers else torch.empty(shape, device=dev, dtype=torch.float32)
            assert tuple(g.shape) == tuple(shape) and g.is_contiguous() and g.dtype == torch.float32, n
            out[n if n in ("d_block_input", "d_input_embed", "pred") else "transformer." + n] = g
            bn = n.encode()
            keep.append(bn)
            slots.append(_lib.Grad(bn, g.data_ptr(), g.numel(), events.get(n) if events else None))
        arr = (_lib.Grad * len(slots))(*slots)
        loss = torch.empty((1,), device=dev, dtype=torch.float32)
        _lib.check(_lib.lib().f5b_cfm_loss_tail_grads(
            tr.plan(), x1.data_ptr(), q["x0"].data_ptr(), q["tx"].data_ptr(), q["t_host"].data_ptr(), q["ln"].data_ptr(),
            q["sp"].data_ptr(), q["batch"], q["seq_len"], q["tx"].shape[1], int(q["drop_audio_cond"]),
            int(q["drop_text"]), loss.data_ptr(), arr, len(slots), _lib.mem_kind(x1), _lib.current_stream_ptr(dev)))
        return loss[0], out

    def prepare_inputs(self, cond, text, duration, *, lens=None, steps=32, sway_sampling_coef=None, seed=None,
                       max_duration=65536, use_epss=True, no_ref_audio=False, duplicate_test=False, t_inter=0.1,
                       edit_mask=None, y0=None) -> dict:
        """Everything `sample` does on the host before the ODE (cfm.py:104-216): integer / mask work that must match
        the reference exactly, the seeded noise and the time grid."""
        if cond.ndim == 2:  # raw wave -> mel (cfm.py:106-109)
            cond = self.mel_spec(cond)
            cond = cond.permute(0, 2, 1)
            assert cond.shape[-1] == self.num_channels
        cond = cond.to(torch.float32)
        batch, cond_seq_len, device = *cond.shape[:2], cond.device
        if not exists(lens):
            lens = torch.full((batch,), cond_seq_len, device=device, dtype=torch.long)

        if isinstance(text, list):  # cfm.py:119-124
            if exists(self.vocab_char_map):
                text = list_str_to_idx(text, self.vocab_char_map).to(device)
            else:
                text = list_str_to_tensor(text).to(device)
            assert text.shape[0] == batch

        cond_mask = lens_to_mask(lens)  # cfm.py:128-130
        if edit_mask is not None:
            cond_mask = cond_mask & edit_mask
        if isinstance(duration, int):
            duration = torch.full((batch,), duration, device=device, dtype=torch.long)
        duration = torch.maximum(torch.maximum((text != -1).sum(dim=-1), lens) + 1, duration)  # cfm.py:135-137
        duration = duration.clamp(max=max_duration)
        max_duration = int(duration.amax())

        if duplicate_test:
            test_cond = F.pad(cond, (0, 0, cond_seq_len, max_duration - 2 * cond_seq_len), value=0.0)
        cond = F.pad(cond, (0, 0, 0, max_duration - cond_seq_len), value=0.0)
        if no_ref_audio:
            cond = torch.zeros_like(cond)
        cond_mask = F.pad(cond_mask, (0, max_duration - cond_mask.shape[-1]), value=False)
        cond_mask = cond_mask.unsqueeze(-1)
        step_cond = torch.where(cond_mask, cond, torch.zeros_like(cond))  # cfm.py:151-153
        has_mask = batch > 1  # cfm.py:155-158: mask = lens_to_mask(duration) if batch > 1 else None

        if y0 is None:  # cfm.py:196-201
            rows = []
            for dur in duration:
                if exists(seed):
                    torch.manual_seed(seed)
                rows.append(torch.randn(int(dur), self.num_channels, device=self.device, dtype=step_cond.dtype))
            y0 = pad_sequence(rows, padding_value=0, batch_first=True)
        y0 = y0.to(device=device, dtype=torch.float32)

        t_start = 0
        if duplicate_test:
            t_start = t_inter
            y0 = (1 - t_start) * y0 + t_start * test_cond
            steps = int(steps * (1 - t_start))
        # the grid is consumed as host floats by the C ABI: built on the CPU (no device launches, no D2H sync)
        t = sway_time_grid(steps, sway_sampling_coef, "cpu", torch.float32, use_epss, t_start)

        return dict(cond=cond, cond_mask=cond_mask, step_cond=step_cond, text=text, duration=duration, y0=y0, t=t,
                    steps=steps, has_mask=has_mask)

    @torch.no_grad()
    def sample(self, cond, text, duration, *, lens=None, steps=32, cfg_strength=1.0, sway_sampling_coef=None,
               seed: int | None = None, max_duration=65536, vocoder: Callable | None = None, use_epss=True,
               no_ref_audio=False, duplicate_test=False, t_inter=0.1, edit_mask=None, y0: torch.Tensor | None = None,
               return_trajectory=True):
        """Same arguments and results as the reference's `CFM.sample` (cfm.py:84-229).  Two keyword-only extras:
        `y0` injects the initial noise (parity tests feed the oracle's CPU noise), `return_trajectory=False` skips
        materialising the [steps+1, b, n, d] trajectory (the reference always builds it)."""
        if self.training:  # (nn.Module.eval() walks all 421 sub-modules: ~1 ms of Python per call when done blindly)
            self.eval()
        p = self.prepare_inputs(cond, text, duration, lens=lens, steps=steps, sway_sampling_coef=sway_sampling_coef,
                                seed=seed, max_duration=max_duration, use_epss=use_epss, no_ref_audio=no_ref_audio,
                                duplicate_test=duplicate_test, t_inter=t_inter, edit_mask=edit_mask, y0=y0)
        cond, cond_mask, step_cond, text, duration = p["cond"], p["cond_mask"], p["step_cond"], p["text"], p["duration"]
        y0, t, steps, has_mask = p["y0"], p["t"], p["steps"], p["has_mask"]

        out_last, trajectory = self._ode(y0, step_cond, text, duration if has_mask else None, t, steps, cfg_strength,
                                         return_trajectory)
        self.transformer.clear_cache()

        out = torch.where(cond_mask, cond, out_last)  # cfm.py:221-223
        if exists(vocoder):
            out = out.permute(0, 2, 1)
            out = vocoder(out)
        return out, trajectory

    def _ode(self, y0, step_cond, text, durations, t, steps, cfg_strength, want_traj):
        """One C-ABI call = the whole `odeint(fn, y0, t, method="euler")` of cfm.py:162-218."""
        plan = self.transformer.plan()
        dev = y0.device
        b, n, mel = y0.shape
        y0 = y0.contiguous()
        sc = step_cond.to(torch.float32).contiguous()
        tx = text.to(device=dev, dtype=torch.long).contiguous()
        dur = None if durations is None else durations.to(device=dev, dtype=torch.int32).contiguous()
        t_host = t.detach().to("cpu", torch.float32).contiguous()
        y_out = torch.empty_like(y0)
        traj = torch.empty((steps + 1, b, n, mel), device=dev, dtype=torch.float32) if want_traj else None
        _lib.check(_lib.lib().f5b_sample(
            plan, y0.data_ptr(), sc.data_ptr(), tx.data_ptr(), _lib.ptr(dur), b, n, tx.shape[1], t_host.data_ptr(),
            steps, float(cfg_strength), y_out.data_ptr(), _lib.ptr(traj), _lib.mem_kind(y0),
            _lib.current_stream_ptr(dev)))
        return y_out, traj
