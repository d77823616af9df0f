"""One data-parallel training step of the flow-matching model (BASELINE.json configs[4]; reference
src/f5_tts/model/trainer.py:59-70 DDP through accelerate, :146 `AdamW(model.parameters(), lr)`, :384 `backward(loss)`,
:387 `clip_grad_norm_(max_grad_norm=1.0)`, :388-390 optimizer / scheduler step, :395 EMA update).

    loss, grads  <- f5b_cfm_loss_tail_grads   forward in the plan's precision mode + backward of every DiT parameter
    all-reduce   <- ONE NCCL all-reduce of the flat gradient buffer over NVLink (sum; the 1 / world average is folded
                    into the optimiser's gradient read) -- the only collective of the path
    clip         <- global L2 norm of the averaged gradient (f5b_op_sumsq), coefficient min(1, max_norm / (norm + 1e-6))
    update       <- f5b_op_adamw over the flat parameter / moment buffers (+ EMA copy)

Parameters, gradients and both Adam moments live in FLAT fp32 device buffers; the module's `nn.Parameter`s are views
into the flat parameter buffer (so `state_dict()` / checkpoints keep the reference's names and shapes) and the backward
pass writes each gradient straight into its slice of the flat gradient buffer.  The learning-rate schedule is the
caller's (pass `lr=` per step), like the reference's scheduler object."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib


class TrainStep:
    def __init__(self, cfm, lr: float = 7.5e-5, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-2,
                 max_grad_norm: float = 1.0, ema_decay: Optional[float] = None, process_group=None):
        self.cfm, self.tr = cfm, cfm.transformer
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.max_grad_norm, self.ema_decay, self.group = max_grad_norm, ema_decay, process_group
        params = [(n, p) for n, p in self.tr.named_parameters()]
        self.names = [n for n, _ in params]
        dev = params[0][1].device
        if dev.type != "cuda":
            raise _lib.F5BError("TrainStep needs the model on a CUDA device (no CPU fallback)")
        self.device = dev
        sizes = [p.numel() for _, p in params]
        self.offsets = [0]
        for s in sizes:
            self.offsets.append(self.offsets[-1] + s)
        n = self.offsets[-1]
        self.flat_param = torch.empty(n, device=dev, dtype=torch.float32)
        self.flat_grad = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(n, device=dev, dtype=torch.float32)
        with torch.no_grad():
            for (_, p), o in zip(params, self.offsets):
                self.flat_param[o:o + p.numel()].copy_(p.detach().reshape(-1))
                p.data = self.flat_param[o:o + p.numel()].view(p.shape)  # the module now reads the flat buffer
        self.ema = self.flat_param.clone() if ema_decay is not None else None
        self.grad_views = {n: self.flat_grad[o:o + s].view(p.shape) for (n, p), o, s in zip(params, self.offsets, sizes)}
        self._scratch = torch.empty(1025, device=dev, dtype=torch.float64)
        self.step_count = 0
        self.last = {}
        self.tr.invalidate_plan()

    @property
    def world(self) -> int:
        import torch.distributed as dist

        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    def reduce_gradients(self):
        """The data-parallel exchange: one all-reduce (sum) of the flat gradient buffer.  Returns the factor that turns
        the sum into the mean (folded into the optimiser's gradient read instead of a separate pass)."""
        import torch.distributed as dist

        w = self.world
        if w > 1:
            dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM, group=self.group)
        return 1.0 / w

    def grad_norm(self, scale: float = 1.0) -> torch.Tensor:
        """L2 norm of scale * flat_grad as a 0-dim device tensor (no host synchronisation)."""
        L, st = _lib.lib(), _lib.current_stream_ptr(self.device)
        _lib.check(L.f5b_op_sumsq(self.flat_grad.data_ptr(), self.flat_grad.numel(), self._scratch.data_ptr(),
                                  self._scratch[1024:].data_ptr(), _lib.device_index(self.device), st))
        return self._scratch[1024].sqrt() * scale

    def step(self, inp, text, *, lens=None, draws: dict | None = None, lr: Optional[float] = None):
        """forward + backward + all-reduce + clip + AdamW (+ EMA).  Returns the loss (0-dim device tensor)."""
        loss, _ = self.cfm.tail_gradients(inp, text, lens=lens, draws=draws, names=self.names, buffers=self.grad_views)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record()
        avg = self.reduce_gradients()
        ev[1].record()
        norm = self.grad_norm(avg)
        coef = 1.0
        if self.max_grad_norm and self.max_grad_norm > 0:  # torch.nn.utils.clip_grad_norm_: min(1, max / (norm + 1e-6))
            coef = min(1.0, self.max_grad_norm / (float(norm) + 1e-6))
        self.step_count += 1
        L = _lib.lib()
        _lib.check(L.f5b_op_adamw(self.flat_param.data_ptr(), self.flat_grad.data_ptr(), self.exp_avg.data_ptr(),
                                  self.exp_avg_sq.data_ptr(), None if self.ema is None else self.ema.data_ptr(),
                                  self.flat_param.numel(), float(self.lr if lr is None else lr), self.betas[0],
                                  self.betas[1], self.eps, self.weight_decay, self.step_count, avg * coef,
                                  float(self.ema_decay or 0.0), _lib.device_index(self.device),
                                  _lib.current_stream_ptr(self.device)))
        ev[2].record()
        self.tr.refresh_plan()  # the plan packs its own 16-bit copies of the weights: re-packed from the new values
        self.last = dict(grad_norm=float(norm), clip_coef=coef, events=ev)
        return loss

    def timings_ms(self) -> dict:
        """all-reduce and optimiser time of the last step (CUDA events on the current stream)."""
        e = self.last["events"]
        e[2].synchronize()
        return {"allreduce_ms": e[0].elapsed_time(e[1]), "clip_adamw_ms": e[1].elapsed_time(e[2])}

    def ema_state_dict(self) -> dict:
        """EMA weights under the reference's names (what trainer.py saves as ema_model_state_dict)."""
        assert self.ema is not None
        out = {}
        for n, o, o2 in zip(self.names, self.offsets[:-1], self.offsets[1:]):
            out["ema_model.transformer." + n] = self.ema[o:o2].view(self.grad_views[n].shape).clone()
        return out
