"""CPU: host-side mirror of the reference interface -- integer / mask construction must match exactly, module trees
must be strict state-dict compatible with the reference's parameter names and shapes."""
import os

import numpy as np
import pytest
import torch

import f5b200
from helpers import O


def test_helpers_match_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "host_helpers.npz"))
    t = f5b200.sway_time_grid(16, -1.0, "cpu", torch.float32)
    assert np.array_equal(t.numpy(), g["t16_sway"])  # bit-exact grid
    assert np.array_equal(f5b200.lens_to_mask(torch.tensor([3, 5, 1])).numpy(), g["mask"])
    idx = f5b200.list_str_to_idx([["a", "b", "?"], ["b"]], {"a": 1, "b": 2})
    assert np.array_equal(idx.numpy(), g["idx"])
    for n in (5, 6, 7, 10, 12, 16, 32, 9):
        assert torch.equal(f5b200.sway_time_grid(n, -1.0, "cpu", torch.float32), O.time_grid(n, -1.0))
        assert torch.equal(f5b200.sway_time_grid(n, None, "cpu", torch.float32, use_epss=False), O.time_grid(n, None, False))


@pytest.mark.parametrize("name", ["F5TTS_v1_Base", "F5TTS_Small", "E2TTS_Base", "F5TTS_Base"])
def test_state_dict_names_and_shapes(name):
    cfg = O.ArchCfg.named(name, depth=2 if name != "E2TTS_Base" else 4)
    cls = f5b200.DiT if cfg.backbone == "DiT" else f5b200.UNetT
    m = cls(dim=cfg.dim, depth=cfg.depth, heads=cfg.heads, ff_mult=cfg.ff_mult, text_dim=cfg.text_dim,
            conv_layers=cfg.conv_layers, text_num_embeds=cfg.text_num_embeds, pe_attn_head=cfg.pe_attn_head,
            text_mask_padding=cfg.text_mask_padding)
    want = O.backbone_param_shapes(cfg)  # pinned against the reference modules by oracle/make_golden.py (strict load)
    got = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    # x_transformers' RotaryEmbedding keeps inv_freq as a persistent buffer: reference checkpoints carry this key
    assert got.pop("rotary_embed.inv_freq") == (32,)
    assert got == {k: tuple(v) for k, v in want.items()}
    if cfg.backbone == "DiT":  # reference zero-initialisation (dit.py:214-224, modules.py:234-235)
        sd = m.state_dict()
        assert float(sd["proj_out.weight"].abs().sum()) == 0 and float(sd["norm_out.linear.bias"].abs().sum()) == 0
        assert float(sd["transformer_blocks.0.attn_norm.linear.weight"].abs().sum()) == 0
        assert float(sd["text_embed.text_blocks.0.grn.gamma"].abs().sum()) == 0


def test_vocos_state_dict():
    v = f5b200.Vocos()
    got = {k: tuple(t.shape) for k, t in v.state_dict().items()}
    assert got == {k: tuple(s) for k, s in O.vocos_param_shapes().items()}
    sd = dict(O.synth_vocos_weights(0))
    sd["head.istft.window"] = torch.hann_window(1024)  # buffers of the released checkpoint are accepted
    sd["feature_extractor.mel_spec.spectrogram.window"] = torch.hann_window(1024)
    v.load_state_dict(sd)


@pytest.mark.parametrize("B", [1, 3])
def test_prepare_inputs_matches_oracle(B):
    cfg = O.ArchCfg(dim=128, depth=1, heads=2, ff_mult=2, text_dim=64, conv_layers=1)
    m = f5b200.DiT(dim=128, depth=1, heads=2, ff_mult=2, text_dim=64, conv_layers=1, text_num_embeds=cfg.text_num_embeds)
    cfm = f5b200.CFM(transformer=m)
    cond, text = O.synth_inputs(40, 30, 12, batch=B, seed=5)
    if B > 1:
        text[1, 7:] = -1
    lens = torch.tensor([30, 22, 17][:B])
    durs = torch.tensor([40, 10, 33][:B])  # second one is lifted to max(text, lens) + 1
    p = cfm.prepare_inputs(cond, text, durs, lens=lens, steps=16, sway_sampling_coef=-1.0, seed=3)
    c2, t2, d2, cm2, sc2, mask2 = O.prepare_sample(cond, text, durs, lens)
    assert torch.equal(p["duration"], d2) and torch.equal(p["cond_mask"], cm2) and torch.equal(p["step_cond"], sc2)
    assert torch.equal(p["cond"], c2) and p["has_mask"] == (mask2 is not None)
    assert torch.equal(p["y0"], O.seeded_noise(d2, 100, 3))  # same seeded CPU noise as the reference
    assert torch.equal(p["t"], O.time_grid(16, -1.0))
    edit = torch.ones(B, 30, dtype=torch.bool)
    edit[:, 5:9] = False
    p = cfm.prepare_inputs(cond, text, 40, lens=lens, edit_mask=edit, no_ref_audio=False, seed=1)
    _, _, _, cm3, sc3, _ = O.prepare_sample(cond, text, 40, lens, edit_mask=edit)
    assert torch.equal(p["cond_mask"], cm3) and torch.equal(p["step_cond"], sc3)


def test_unsupported_options_raise():
    with pytest.raises(ValueError):
        f5b200.DiT(dim=128, qk_norm="rms_norm")
    with pytest.raises(ValueError):
        f5b200.MelSpec(mel_spec_type="bigvgan")
    with pytest.raises(ValueError):
        f5b200.CFM(transformer=f5b200.DiT(dim=128, depth=1, heads=2), odeint_kwargs=dict(method="midpoint"))
    if not torch.cuda.is_available():  # the loss path has no CPU fallback either: it must fail loudly, not fall back
        with pytest.raises(RuntimeError):
            f5b200.CFM(transformer=f5b200.DiT(dim=128, depth=1, heads=2))(torch.zeros(1, 4, 100), torch.zeros(1, 2).long())


def test_precision_switches_keep_fp32_masters():
    m = f5b200.DiT(dim=128, depth=1, heads=2)
    assert m.half().precision == "bf16" and m.proj_out.weight.dtype == torch.float32
    assert m.to(torch.float32).precision == "fp32"
    assert m.to(dtype=torch.float16).precision == "bf16" and m.proj_out.weight.dtype == torch.float32


def test_sharding_partition():
    from f5b200.sharding import micro_batches, request_cost, shard_requests

    g = torch.Generator().manual_seed(0)
    lengths = torch.randint(281, 1876, (256,), generator=g).tolist()
    for ws in (1, 2, 4, 8):
        shards = shard_requests(lengths, ws)
        assert sorted(i for s in shards for i in s) == list(range(256))
        loads = [sum(request_cost(lengths[i]) for i in s) for s in shards]
        assert max(loads) / (sum(loads) / ws) < 1.02  # LPT keeps the imbalance under 2 %
        for s in shards:
            mb = micro_batches(s, lengths, max_rows=16384)
            assert sorted(i for b in mb for i in b) == sorted(s)
            assert all(len(b) * max(lengths[i] for i in b) <= 16384 for b in mb)
