"""Request sharding for multi-GPU inference: utterances are independent, so each GPU takes a disjoint shard and no
collective runs on the data path (the reference shards the same way with `accelerator.split_between_processes`,
src/f5_tts/eval/eval_infer_batch.py:181, and `DistributedSampler`, runtime/triton_trtllm/benchmark.py:340-341).

Unlike the reference's contiguous split, work is balanced by an attention-aware cost.  A job whose whole request list
is known (`plan_job`) is cut into length-sorted micro-batches first and the batches are dealt to the ranks; a rank that
only knows its own utterances (`my_shard`) cuts its shard into micro-batches itself."""
from __future__ import annotations

from typing import List, Sequence


def request_cost(n_frames: int, dim: int = 1024, depth: int = 22, ff_mult: int = 2) -> float:
    """Algorithmic flops of one CFG-doubled backbone evaluation over n_frames: per token per layer
    (8 + 4 ff_mult) d^2 for the projections and 4 n d for attention; times two branches."""
    per_token = depth * ((8 + 4 * ff_mult) * dim * dim + 4 * n_frames * dim)
    return 2.0 * n_frames * per_token


def shard_requests(lengths: Sequence[int], world_size: int, **cost_kw) -> List[List[int]]:
    """Longest-processing-time-first assignment.  Deterministic; every index appears in exactly one shard."""
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    loads = [0.0] * world_size
    shards: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (loads[k], k))
        shards[r].append(i)
        loads[r] += request_cost(lengths[i], **cost_kw)
    return shards


def micro_batches(indices: Sequence[int], lengths: Sequence[int], max_rows: int, max_batch: int = 64,
                  max_waste: float = 0.05) -> List[List[int]]:
    """Cut a shard into batches of similar length: sorted by length, a batch closes when batch_size * longest would
    exceed max_rows (padded rows) or max_batch utterances, or when admitting the next (shorter) utterance would push the
    batch's padding -- 1 - sum(lengths) / (batch_size * longest) -- over `max_waste`.  The last rule is what keeps a
    small shard efficient: with 32 utterances per GPU (C3 on 8 GPUs) row-count-only batches padded 24 % of their frames,
    waste-bounded ones 4 % (at 5 %: 3.9-4.6 % for 1 / 2 / 4 / 8 GPUs)."""
    out: List[List[int]] = []
    cur: List[int] = []
    total = 0
    for i in sorted(indices, key=lambda j: (-lengths[j], j)):
        if cur:
            longest = lengths[cur[0]]
            n = len(cur) + 1
            waste = 1.0 - (total + lengths[i]) / (n * longest)
            if n * longest > max_rows or len(cur) >= max_batch or waste > max_waste:
                out.append(cur)
                cur, total = [], 0
        cur.append(i)
        total += lengths[i]
    if cur:
        out.append(cur)
    return out


def batch_cost(group: Sequence[int], lengths: Sequence[int], **cost_kw) -> float:
    """Cost of one micro-batch: every row is computed at the padded length."""
    return len(group) * request_cost(max(lengths[i] for i in group), **cost_kw)


def plan_job(lengths: Sequence[int], world_size: int, max_rows: int = 16384, max_batch: int = 64,
             max_waste: float = 0.05, **cost_kw):
    """BATCH FIRST, SHARD SECOND: the whole request list is cut into length-sorted micro-batches (`micro_batches`), then
    the BATCHES are dealt to the ranks longest-processing-time first.  Sharding utterances first (`shard_requests` +
    `micro_batches` per rank) leaves every rank with a thin slice of each length class -- on 8 GPUs the C3 job ran as
    12-14 calls of ~2.5 utterances per rank, at 13 % lower tensor throughput than the 16 k-row batches of one GPU; dealt
    as whole batches the ranks keep the single-GPU batch sizes and padding.  With few batches per rank the balance depends
    on the batch size, so a few row limits are tried and the one with the smallest estimated finishing time is kept (a
    smaller limit is charged 3 % per halving for its lower kernel efficiency).
    Returns (per-rank list of micro-batches, the row limit used).  Deterministic: every rank computes the same plan."""
    import math

    best = None
    for rows in sorted({max_rows, max_rows * 7 // 8, max_rows * 3 // 4, max_rows * 5 // 8, max_rows // 2}, reverse=True):
        groups = micro_batches(range(len(lengths)), lengths, rows, max_batch, max_waste)
        order = sorted(range(len(groups)), key=lambda k: (-batch_cost(groups[k], lengths, **cost_kw), k))
        loads = [0.0] * world_size
        plan: List[List[List[int]]] = [[] for _ in range(world_size)]
        for k in order:
            r = min(range(world_size), key=lambda q: (loads[q], q))
            plan[r].append(groups[k])
            loads[r] += batch_cost(groups[k], lengths, **cost_kw)
        est = max(loads) / (1.0 - 0.03 * math.log2(max_rows / rows))
        if best is None or est < best[0] * (1.0 - 1e-9):
            best = (est, plan, rows)
    return best[1], best[2]


def my_shard(lengths: Sequence[int], rank: int, world_size: int, **cost_kw) -> List[int]:
    """The utterance indices this rank synthesises (every rank computes the same deterministic partition)."""
    return shard_requests(lengths, world_size, **cost_kw)[rank]


def gather_to_rank0(local: dict, group=None) -> dict | None:
    """Host-side gather of `{utterance index: result}` dicts (the reference gathers on the host as well,
    eval_infer_batch.py:181-208).  No tensor collective is involved: the data path stays shard-local."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return dict(local)
    parts = [None] * dist.get_world_size(group) if dist.get_rank(group) == 0 else None
    dist.gather_object(local, parts, dst=0, group=group)
    if parts is None:
        return None
    merged: dict = {}
    for p in parts:
        overlap = merged.keys() & p.keys()
        assert not overlap, f"utterances synthesised twice: {sorted(overlap)}"
        merged.update(p)
    return merged


def max_over_ranks(seconds: float, device=None, group=None) -> float:
    """Timing rule of bench.py: a multi-GPU number is the MAX over ranks of the device-timed region."""
    import torch
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(seconds)
    t = torch.tensor([seconds], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
