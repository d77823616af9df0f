"""N > 1 host logic on CPU: two gloo ranks shard a request list, "synthesise" their shard, gather on rank 0 and take the
max-over-ranks timing -- the same plumbing bench.py runs over NCCL on the GPU box (no data-path collective)."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "korean-f5-tts_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _worker(rank, world, port, lengths, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from f5b200 import sharding

        mine = sharding.my_shard(lengths, rank, world)
        local = {i: lengths[i] * 256 for i in mine}  # stand-in for a waveform: its sample count
        merged = sharding.gather_to_rank0(local)
        t = sharding.max_over_ranks(0.5 + rank)  # rank 1 is the slow one
        dist.barrier()
        if rank == 0:
            out_q.put((merged, t, mine))
    finally:
        dist.destroy_process_group()


def test_two_rank_request_sharding_gloo():
    g = torch.Generator().manual_seed(0)
    lengths = [int(x) for x in torch.randint(281, 1875, (37,), generator=g)]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lengths, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged, t, mine0 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(merged) == list(range(len(lengths)))  # every utterance exactly once
    assert all(merged[i] == lengths[i] * 256 for i in merged)
    assert t == 1.5  # max over ranks
    from f5b200 import sharding

    loads = [sum(sharding.request_cost(lengths[i]) for i in s) for s in sharding.shard_requests(lengths, 2)]
    assert max(loads) / min(loads) < 1.05  # LPT balance
    assert mine0 == sharding.shard_requests(lengths, 2)[0]


def _ddp_worker(rank, world, port, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from f5b200.train import TrainStep

        ts = TrainStep.__new__(TrainStep)  # the exchange itself: no device, no model needed
        ts.group = None
        ts.flat_grad = torch.arange(10, dtype=torch.float32) * (rank + 1)  # rank r holds (r + 1) * [0..9]
        ts.buckets = [(6, 10), (2, 6), (0, 2)]  # exchanged bucket by bucket, in the backward pass's order
        avg = ts.reduce_gradients()
        dist.barrier()
        out_q.put((rank, avg, ts.flat_grad.clone()))
    finally:
        dist.destroy_process_group()


def test_bucket_ranges_follow_the_backward_pass():
    from f5b200.train import bucket_ranges

    names = ["time_embed.w", "text_embed.w", "input_embed.w", "transformer_blocks.0.a", "transformer_blocks.0.b",
             "transformer_blocks.1.a", "transformer_blocks.1.b", "transformer_blocks.2.a", "norm_out.w", "proj_out.w"]
    offsets = [0, 3, 5, 9, 10, 14, 15, 20, 22, 25, 31]
    # [block 2 + norm_out + proj_out], block 1, block 0, [embeddings]: every element exactly once
    assert bucket_ranges(names, offsets, 31) == [(20, 31), (14, 20), (9, 14), (0, 9)]


def test_two_rank_gradient_exchange_gloo():
    """The training step's only collective (reference trainer.py:59-70: DDP averages gradients): all-reduce (sum) of the
    flat gradient buffer, bucket by bucket; the 1 / world factor is returned and folded into the optimiser's gradient
    read."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, avg, g in res:
        assert avg == 0.5 and torch.equal(g, torch.arange(10, dtype=torch.float32) * 3)  # sum on every rank
        assert torch.equal(g * avg, torch.arange(10, dtype=torch.float32) * 1.5)          # the mean of the two ranks


def test_plan_job_deals_whole_batches():
    """Batch first, shard second: every utterance lands in exactly one batch of exactly one rank, batches respect the
    row limit, the plan is deterministic, and on a C3-like length mix the estimated finishing time of the slowest of 8
    ranks stays within 8 % of the ideal split with single-GPU batch sizes (>= 3 utterances per batch on average)."""
    import random

    from f5b200 import sharding

    rnd = random.Random(5)
    lengths = [rnd.randint(282, 1874) for _ in range(256)]
    for world in (1, 2, 4, 8):
        plan, rows = sharding.plan_job(lengths, world, max_rows=16384)
        assert plan == sharding.plan_job(lengths, world, max_rows=16384)[0]
        seen = sorted(i for r in plan for g in r for i in g)
        assert seen == list(range(256))
        assert all(len(g) * max(lengths[i] for i in g) <= rows or len(g) == 1 for r in plan for g in r)
        loads = [sum(sharding.batch_cost(g, lengths) for g in r) for r in plan]
        ideal = sum(sharding.request_cost(n) for n in lengths) / world
        assert max(loads) <= 1.08 * ideal, (world, max(loads) / ideal)  # padding (<= 5 %) + balance
        assert sum(len(r) for r in plan) <= 256 / 3
