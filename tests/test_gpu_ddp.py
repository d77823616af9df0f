"""Data-parallel training step on TWO GPUs over NCCL (reference trainer.py:59-70: DDP averages the gradients of the
ranks' micro-batches): after one `TrainStep.step` on different data the two ranks hold identical parameters, the
bucketed exchange that overlaps the backward pass gives the bits of the plain one-buffer all-reduce, and the result is
the single-process update driven by the mean of the two ranks' gradients.  Skipped on a one-GPU box."""
import os

import pytest
import torch
import torch.multiprocessing as mp

from helpers import O, build_cfm

pytestmark = pytest.mark.gpu


def _draws(d):
    return dict(frac_lengths=d.frac_lengths, rand=d.rand, x0=d.x0, time=d.time, drop_audio_cond=d.drop_audio_cond,
                drop_text=d.drop_text)


def _case(rank):
    x1, text = O.synth_inputs(60, 60, 12, batch=2, seed=3 + rank)
    return x1, text, torch.tensor([60, 44 + 8 * rank]), O.draw_loss_randoms(x1, 70 + rank)


def _worker(rank, world, port, out_q):
    import torch.distributed as dist

    import f5b200

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        cfg = O.ArchCfg(dim=128, depth=3, heads=2, ff_mult=2, text_dim=64, conv_layers=2, attn_mask_enabled=True)
        sd = O.synth_backbone_weights(cfg, seed=5)
        x1, text, lens, d = _case(rank)
        dev = torch.device("cuda", rank)
        res = {}
        for mode in ("overlapped", "plain"):
            cfm = build_cfm(cfg, sd, "fp32", device=dev)
            ts = f5b200.TrainStep(cfm, lr=1e-3, weight_decay=1e-2, max_grad_norm=0.05)
            if mode == "plain":
                ts.reduce_gradients = lambda ts=ts: f5b200.TrainStep.reduce_gradients(ts, overlapped=False)
            with torch.cuda.stream(torch.cuda.Stream(device=dev)):
                ts.step(x1.to(dev), text.to(dev), lens=lens.to(dev), draws=_draws(d))
                torch.cuda.current_stream().synchronize()
            res[mode] = (ts.flat_param.detach().cpu(), ts.last["grad_norm"], len(ts.buckets))
        out_q.put((rank, res))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_step_equals_the_averaged_single_process_step():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    p0, norm0, nb = res[0]["overlapped"]
    assert nb == 4  # [block 2 + norm_out + proj_out], block 1, block 0, embeddings
    for rank in (0, 1):
        for mode in ("overlapped", "plain"):
            assert torch.equal(res[rank][mode][0], p0), (rank, mode)  # same bits on both ranks, bucketed == plain
            assert res[rank][mode][1] == norm0

    # single process: AdamW on the MEAN of the two ranks' oracle gradients, clipped by the norm of that mean
    cfg = O.ArchCfg(dim=128, depth=3, heads=2, ff_mult=2, text_dim=64, conv_layers=2, attn_mask_enabled=True)
    sd = O.synth_backbone_weights(cfg, seed=5)
    leaves = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    total = 0.0
    for rank in (0, 1):
        x1, text, lens, d = _case(rank)
        loss, _, _ = O.cfm_loss(leaves, cfg, x1, text, lens, d)
        total = total + 0.5 * loss
    total.backward()
    opt = torch.optim.AdamW(list(leaves.values()), lr=1e-3, weight_decay=1e-2)
    norm = torch.nn.utils.clip_grad_norm_(list(leaves.values()), 0.05)
    opt.step()
    assert abs(norm0 - float(norm)) <= 1e-4 * float(norm)
    cfm = build_cfm(cfg, sd, "fp32")
    import f5b200

    ts = f5b200.TrainStep(cfm, lr=1e-3)  # only for the name -> flat-buffer offsets
    worst = 0.0
    for n, a, b in zip(ts.names, ts.offsets[:-1], ts.offsets[1:]):
        ref = leaves[n].detach().reshape(-1)
        worst = max(worst, float((p0[a:b] - ref).abs().max()) / max(float(ref.abs().max()), 1e-8))
    print(f"two-rank step: parameters within {worst:.2e} (relative to max|w|) of the averaged single-process AdamW step")
    assert worst <= 1e-4
