"""One data-parallel training step of the flow-matching model (BASELINE.json configs[4]; reference
src/f5_tts/model/trainer.py:59-70 DDP through accelerate, :146 `AdamW(model.parameters(), lr)`, :384 `backward(loss)`,
:387 `clip_grad_norm_(max_grad_norm=1.0)`, :388-390 optimizer / scheduler step, :395 EMA update).

    loss, grads  <- f5b_cfm_loss_tail_grads   forward in the plan's precision mode + backward of every DiT parameter
    all-reduce   <- NCCL all-reduce (sum) of the flat gradient buffer over NVLink, in BUCKETS that follow the backward
                    pass: [last block .. proj_out], then each earlier block, then the embeddings.  The library records a
                    CUDA event per bucket as its last gradient lands; a communication stream waits for it, so the
                    exchange of block l runs under the backward of blocks l-1 .. 0 (what DDP's bucketing does for the
                    reference, trainer.py:59-70).  The 1 / world average is folded into the optimiser's gradient read
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


def bucket_ranges(names, offsets, total):
    """All-reduce buckets over the flat gradient buffer in the order the backward pass completes them: [last block ..
    proj_out], block depth-2, ..., block 0, [embeddings].  The ranges are contiguous because named_parameters() lists
    time_embed, text_embed, input_embed, transformer_blocks.0 .. depth-1, norm_out, proj_out in that order."""
    first = {}
    for nme, o in zip(names, offsets):
        if nme.startswith("transformer_blocks."):
            key = int(nme.split(".")[1])
        else:  # the embeddings come before block 0; norm_out / proj_out follow the last block and complete with it
            key = max(first) if any(k >= 0 for k in first) else -1
        first.setdefault(key, o)
    keys = sorted(first)
    bounds = [first[k] for k in keys] + [total]
    assert bounds == sorted(bounds), "parameters of one block are not contiguous in the flat buffer"
    return [(bounds[i], bounds[i + 1]) for i in range(len(keys))][::-1]


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
        self.buckets = bucket_ranges(self.names, self.offsets, n)  # last block first, embeddings last
        self._bucket_of = {}
        for nme, o in zip(self.names, self.offsets):
            self._bucket_of[nme] = next(i for i, (a, b) in enumerate(self.buckets) if a <= o < b)
        self._events = [torch.cuda.Event() for _ in self.buckets]
        for e in self._events:
            e.record()  # creates the CUDA event (torch makes it lazily) so that its handle can be given to the library
        self._event_of = {nme: self._events[i].cuda_event for nme, i in self._bucket_of.items()}
        self._comm = torch.cuda.Stream(device=dev)
        self.step_count = 0
        self.last = {}
        self.tr.invalidate_plan()

    @property
    def world(self) -> int:
        import torch.distributed as dist

        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    def reduce_gradients(self, overlapped: bool = True):
        """The data-parallel exchange (sum) of the flat gradient buffer.  Called right after the backward pass has been
        ENQUEUED (the library call is asynchronous): every bucket's all-reduce is issued on a communication stream behind
        the event the library records when the bucket's last gradient has been written, so it runs while earlier blocks
        are still being differentiated; the current stream then waits for all of them.  Returns the factor that turns
        the sum into the mean (folded into the optimiser's gradient read instead of a separate pass)."""
        import torch.distributed as dist

        w = self.world
        if w > 1 and (not overlapped or not self.flat_grad.is_cuda):  # bucket after bucket on the current stream
            for a, b in getattr(self, "buckets", None) or [(0, self.flat_grad.numel())]:
                dist.all_reduce(self.flat_grad[a:b], op=dist.ReduceOp.SUM, group=self.group)
        elif w > 1:
            works = []
            for ev, (a, b) in zip(self._events, self.buckets):
                with torch.cuda.stream(self._comm):
                    self._comm.wait_event(ev)
                    works.append(dist.all_reduce(self.flat_grad[a:b], op=dist.ReduceOp.SUM, group=self.group, async_op=True))
            for wk in works:
                wk.wait()  # the current stream waits for the communication (no host synchronisation)
        return 1.0 / w

    def grad_norm(self, scale: float = 1.0) -> torch.Tensor:
        """L2 norm of scale * flat_grad as a 0-dim device tensor (no host synchronisation)."""
        L, st = _lib.lib(), _lib.current_stream_ptr(self.device)
        _lib.check(L.f5b_op_sumsq(self.flat_grad.data_ptr(), self.flat_grad.numel(), self._scratch.data_ptr(),
                                  self._scratch[1024:].data_ptr(), _lib.device_index(self.device), st))
        return self._scratch[1024].sqrt() * scale

    def step(self, inp, text, *, lens=None, draws: dict | None = None, lr: Optional[float] = None):
        """forward + backward + all-reduce + clip + AdamW (+ EMA).  Returns the loss (0-dim device tensor)."""
        loss, _ = self.cfm.tail_gradients(inp, text, lens=lens, draws=draws, names=self.names, buffers=self.grad_views,
                                          events=self._event_of)
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
