"""Golden vectors from the reference's OWN second implementation of the sampling loop -- its TensorRT-LLM runtime
(src/f5_tts/runtime/triton_trtllm) -- for the pieces of the path whose first implementation lives in third-party
packages that are absent from the reference tree (x_transformers' rotary embedding, torchdiffeq's Euler step):

  * model_repo_f5_tts/f5_tts/1/f5_tts_trtllm.py:228-260 (F5TTS.__init__): the RoPE angle table (inv_freq, outer product,
    pairwise repeat, cos / sin), the EPSS tables, the sway-warped time grid and its steps, the sinusoidal time embedding.
  * f5_tts_trtllm.py:362-372 (F5TTS.forward): the classifier-free-guidance combine and the Euler update.
  * patch/f5tts/modules.py:210-276: `rotate_every_two_3dim`, `apply_rotary_pos_emb_3dim` -- HOW the table is applied
    (interleaved pairs, per head of 64, the first `pe_attn_head` heads only).

The files import tensorrt / tensorrt_llm (absent here), so nothing is imported: the statements / functions are lifted
out with `ast` and executed unmodified.  The modules.py functions are written against tensorrt_llm.functional; the few
operators they use (shape, concat, slice, expand_dims, constant, view) are supplied as one-line torch equivalents of
their documented semantics, and a tensor subclass provides TensorRT's `ndim()` method.

Run in the build container (needs /root/reference):  python oracle/make_golden_trt.py  -> tests/golden/trt_witness.npz
"""
import ast
import json
import math
import os
import types

import numpy as np
import torch

RT = "/root/reference/src/f5_tts/runtime/triton_trtllm"
RUNNER = os.path.join(RT, "model_repo_f5_tts/f5_tts/1/f5_tts_trtllm.py")
MODULES = os.path.join(RT, "patch/f5tts/modules.py")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _comp(stmts, path):
    return compile(ast.fix_missing_locations(ast.Module(body=list(stmts), type_ignores=[])), path, "exec")


def _cls_fn(tree, cls, fn):
    c = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls)
    return next(n for n in c.body if isinstance(n, ast.FunctionDef) and n.name == fn)


def _assigns_self(stmt, attr):
    return isinstance(stmt, ast.Assign) and any(
        isinstance(t, ast.Attribute) and isinstance(t.value, ast.Name) and t.value.id == "self" and t.attr == attr
        for t in stmt.targets)


def runner_tables(head_dim=64, max_mel_len=4096):
    """F5TTS.__init__ from `self.n_mel_channels = ...` to its end, executed as written."""
    tree = ast.parse(open(RUNNER, encoding="utf-8").read())
    init = _cls_fn(tree, "F5TTS", "__init__")
    start = next(i for i, s in enumerate(init.body) if _assigns_self(s, "n_mel_channels"))
    me = types.SimpleNamespace(max_mel_len=max_mel_len, device="cpu")
    ns = dict(self=me, config={"pretrained_config": {"mel_dim": 100, "dim_head": head_dim}}, torch=torch, math=math)
    exec(_comp(init.body[start:], RUNNER), ns)
    return me, ns["epss"], ns["time_step"], ns["delta_t"]


def runner_euler(noise_half, pred_cond, pred_uncond, cfg_strength, t_scale):
    """The two statements of F5TTS.forward's loop that combine the guided prediction and advance the state."""
    tree = ast.parse(open(RUNNER, encoding="utf-8").read())
    fwd = _cls_fn(tree, "F5TTS", "forward")
    loop = next(n for n in fwd.body if isinstance(n, ast.For))
    want = []
    for s in loop.body:
        if isinstance(s, ast.Assign) and isinstance(s.targets[0], ast.Name) and s.targets[0].id in ("guidance", "noise_half"):
            want.append(s)
    assert [s.targets[0].id for s in want] == ["guidance", "noise_half"]
    ns = dict(noise_half=noise_half, pred_cond=pred_cond, pred_uncond=pred_uncond, cfg_strength=cfg_strength,
              t_scale=t_scale)
    exec(_comp(want, RUNNER), ns)
    return ns["noise_half"]


class _T(torch.Tensor):
    """torch.Tensor with TensorRT's `ndim()` METHOD (tensorrt_llm.functional.Tensor.ndim)."""

    def ndim(self):  # noqa: D401
        return self.dim()


def _trt_functional():
    """tensorrt_llm.functional operators used by the two lifted functions, by their documented semantics."""
    def shape(t, i):
        return t.shape[i]

    def concat(items, dim=0):
        if all(not torch.is_tensor(v) for v in items):  # a shape tensor assembled from scalars
            return [int(v) for v in items]
        return torch.cat(list(items), dim=dim)

    def slice_(t, starts, sizes, strides=None):
        strides = strides or [1] * len(starts)
        idx = tuple(slice(int(s), int(s) + int(n) * int(k), int(k)) for s, n, k in zip(starts, sizes, strides))
        return t[idx]

    def expand_dims(t, dim):
        return t.unsqueeze(dim)

    def constant(arr):
        return torch.from_numpy(np.asarray(arr))

    def view(t, shp):
        return t.reshape([int(v) for v in shp])

    net = types.SimpleNamespace(plugin_config=types.SimpleNamespace(remove_input_padding=False))
    return dict(shape=shape, concat=concat, slice=slice_, expand_dims=expand_dims, constant=constant, view=view,
                default_net=lambda: net, trt_dtype_to_np=lambda dt: np.float32, np=np, Tensor=torch.Tensor)


def modules_rope():
    tree = ast.parse(open(MODULES, encoding="utf-8").read())
    fns = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("rotate_every_two_3dim",
                                                                               "apply_rotary_pos_emb_3dim")]
    assert len(fns) == 2
    ns = _trt_functional()
    exec(_comp(fns, MODULES), ns)
    return ns["apply_rotary_pos_emb_3dim"]


def main():
    me, epss, time_step, delta_t = runner_tables()
    n = 320
    g = torch.Generator().manual_seed(0)
    out = dict(freqs=me.freqs[0, :n].numpy(), rope_cos_half=me.rope_cos[0, :n].float().numpy(),
               rope_sin_half=me.rope_sin[0, :n].float().numpy(), nfe=np.int64(me.nfe_steps),
               time_step=time_step.numpy(), delta_t=delta_t.numpy(), time_expand=me.time_expand[0].numpy(),
               epss_json=np.frombuffer(json.dumps({str(k): v for k, v in epss.items()}).encode(), dtype=np.uint8))
    # Euler + CFG
    y = torch.randn(2, 50, 100, generator=g)
    pc, pu = torch.randn(2, 50, 100, generator=g), torch.randn(2, 50, 100, generator=g)
    dt = delta_t[3].unsqueeze(0)
    out.update(eu_y=y.numpy(), eu_pc=pc.numpy(), eu_pu=pu.numpy(), eu_dt=dt.numpy(),
               eu_out=runner_euler(y, pc, pu, 2.0, dt).numpy())
    # RoPE application: fp32 cos / sin of the runner's own angle table (the runner rounds them to half for its engine)
    apply = modules_rope()
    heads = 4
    x = torch.randn(2, n, heads * 64, generator=g)
    cos, sin = me.freqs[:, :n].cos(), me.freqs[:, :n].sin()
    out["rope_x"] = x.numpy()
    for name, pe in (("all", None), ("one", 1)):
        r = apply(x.as_subclass(_T), cos.as_subclass(_T), sin.as_subclass(_T), pe)
        out[f"rope_out_{name}"] = torch.Tensor(r).numpy() if not isinstance(r, np.ndarray) else r
    path = os.path.join(ROOT, "tests", "golden", "trt_witness.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
