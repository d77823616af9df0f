"""Kernel-level parity through the C ABI (f5b_op_*): every contraction kernel against a plain PyTorch fp32 reference
of the same op (computed on the host).  Tolerances: fp32 kernels 1e-4 relative to the output scale; bf16 tensor-core
kernels 2e-2 relative (bf16 operand rounding, fp32 accumulation)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from f5b200 import _lib as L

    return L


def _gemm(mode, a, w, bias, act="none", force_bn=0):
    L = _lib()
    dev = torch.device("cuda", 0)
    a_d, w_d = a.to(dev).contiguous(), w.to(dev).contiguous()
    b_d = bias.to(dev).contiguous() if bias is not None else None
    out = torch.empty((a.shape[0], w.shape[0]), device=dev)
    L.check(L.lib().f5b_op_gemm(mode, a_d.data_ptr(), w_d.data_ptr(), L.ptr(b_d), a.shape[0], w.shape[0], a.shape[1],
                                L.ACT[act], out.data_ptr(), force_bn, 0, 0))
    return out.cpu()


def _ref_act(x, act):
    import torch.nn.functional as F

    return {"none": lambda t: t, "gelu_tanh": lambda t: F.gelu(t, approximate="tanh"), "gelu": F.gelu, "mish": F.mish,
            "silu": F.silu}[act](x)


@pytest.mark.parametrize("M,N,K,act", [(300, 256, 128, "none"), (1880, 1024, 712, "gelu_tanh"), (16, 4100, 1024, "silu"),
                                       (33, 100, 300, "mish"), (129, 132, 64, "gelu")])
def test_gemm_fp32(M, N, K, act):
    g = torch.Generator().manual_seed(M + N)
    a = torch.randn(M, K, generator=g)
    w = torch.randn(N, K, generator=g) * K ** -0.5
    b = torch.randn(N, generator=g) * 0.1
    out = _gemm(0, a, w, b, act)
    ref = _ref_act(a.double() @ w.double().T + b.double(), act).float()
    assert (out - ref).abs().max() <= 1e-4 * max(1.0, float(ref.abs().max()))


@pytest.mark.parametrize("M,N,K,act,bn", [(128, 128, 64, "none", 128), (256, 256, 256, "none", 64),
                                          (1880, 3072, 1024, "none", 0), (1880, 1024, 2048, "gelu_tanh", 256),
                                          (1880, 100, 1024, "none", 64), (940, 1024, 712, "none", 128),
                                          (77, 520, 1000, "gelu", 0), (1880, 3072, 1024, "none", 192),
                                          (300, 500, 320, "gelu_tanh", 192), (1880, 2048, 1024, "gelu_tanh", 128),
                                          # CTA pairs (cta_group::2): 1000 + tile width; odd row-tile counts leave half a pair empty
                                          (256, 256, 128, "none", 1128), (1880, 3072, 1024, "none", 1192),
                                          (1880, 2048, 1024, "gelu_tanh", 1256), (1880, 1024, 2048, "none", 1128),
                                          (130, 200, 1000, "gelu", 1256), (940, 1024, 712, "silu", 1192)])
def test_gemm_bf16_tcgen05(M, N, K, act, bn):
    g = torch.Generator().manual_seed(M + N + K)
    a = torch.randn(M, K, generator=g)
    w = torch.randn(N, K, generator=g) * K ** -0.5
    b = torch.randn(N, generator=g) * 0.1
    out = _gemm(1, a, w, b, act, bn)
    ar, wr = a.bfloat16().double(), w.bfloat16().double()  # the kernel rounds operands to bf16, accumulates in fp32
    ref = _ref_act(ar @ wr.T + b.double(), act).float()
    err = (out - ref).abs().max()
    assert err <= 2e-3 * max(1.0, float(ref.abs().max())), f"max err {err}"


def _attention(mode, qkv, kv_len, batches, rows, heads):
    L = _lib()
    dev = torch.device("cuda", 0)
    q_d = qkv.to(dev).contiguous()
    l_d = kv_len.to(dev, torch.int32).contiguous() if kv_len is not None else None
    out = torch.empty((batches * rows, heads * 64), device=dev)
    L.check(L.lib().f5b_op_attention(mode, q_d.data_ptr(), L.ptr(l_d), batches, rows, heads, out.data_ptr(), 0, 0))
    return out.cpu()


def _attention_ref(qkv, kv_len, batches, rows, heads):
    inner = heads * 64
    x = qkv.double().view(batches, rows, 3, heads, 64)
    q, k, v = (x[:, :, i].transpose(1, 2) for i in range(3))
    s = q @ k.transpose(-1, -2) * 0.125
    if kv_len is not None:
        keep = torch.arange(rows)[None, :] < kv_len[:, None]
        s = s.masked_fill(~keep[:, None, None, :], float("-inf"))
    o = torch.softmax(s, -1) @ v
    return o.transpose(1, 2).reshape(batches * rows, inner).float()


@pytest.mark.parametrize("batches,rows,heads,lens", [(1, 100, 2, None), (2, 333, 3, [333, 201]), (2, 940, 16, None)])
def test_attention_fp32(batches, rows, heads, lens):
    g = torch.Generator().manual_seed(rows)
    qkv = torch.randn(batches * rows, 3 * heads * 64, generator=g)
    kv = torch.tensor(lens) if lens else None
    out = _attention(0, qkv, kv, batches, rows, heads)
    ref = _attention_ref(qkv, kv, batches, rows, heads)
    assert (out - ref).abs().max() <= 1e-4


@pytest.mark.parametrize("batches,rows,heads,lens", [(1, 128, 1, None), (1, 100, 2, None), (2, 333, 3, [333, 201]),
                                                     (2, 940, 16, None), (1, 2801, 4, None)])
def test_attention_bf16_tcgen05(batches, rows, heads, lens):
    g = torch.Generator().manual_seed(rows + 1)
    qkv = torch.randn(batches * rows, 3 * heads * 64, generator=g)
    kv = torch.tensor(lens) if lens else None
    out = _attention(1, qkv, kv, batches, rows, heads)
    ref = _attention_ref(qkv.bfloat16().float(), kv, batches, rows, heads)
    err = (out - ref).abs().max()
    assert err <= 2e-2, f"max err {err}"  # P and the output are rounded to bf16


@pytest.mark.parametrize("rows,heads", [(700, 2), (1300, 1)])
def test_attention_bf16_rising_scores_rescale_path(rows, heads):
    """Keys whose scores keep growing along the sequence: every later key block raises the row maximum by far more
    than the lazy-rescaling threshold (2^8), so the O / row-sum rescale in tensor memory is exercised in every block."""
    g = torch.Generator().manual_seed(rows)
    qkv = torch.randn(rows, 3 * heads * 64, generator=g)
    inner = heads * 64
    q = qkv[:, :inner].view(rows, heads, 64)
    k = qkv[:, inner:2 * inner].view(rows, heads, 64)
    # align k with the mean query direction and let its length grow with the position: scores rise by ~60 per block
    d = torch.nn.functional.normalize(q.mean(0), dim=-1)
    q += 3.0 * d[None]
    k += (torch.arange(rows)[:, None, None] / 128.0) * 20.0 * d[None]
    out = _attention(1, qkv, None, 1, rows, heads)
    ref = _attention_ref(qkv.bfloat16().float(), None, 1, rows, heads)
    err = (out - ref).abs().max()
    assert torch.isfinite(out).all() and err <= 3e-2, f"max err {err}"


def test_op_time_runs():
    L = _lib()
    ms = C.c_float()
    L.check(L.lib().f5b_op_time(0, 512, 512, 512, 3, C.byref(ms), 0, 0))
    assert ms.value > 0


def _attention_bwd(mode, qkv, d_out, kv_len, batches, rows, heads):
    L = _lib()
    dev = torch.device("cuda", 0)
    q_d, g_d = qkv.to(dev).contiguous(), d_out.to(dev).contiguous()
    l_d = kv_len.to(dev, torch.int32).contiguous() if kv_len is not None else None
    out = torch.full((batches * rows, 3 * heads * 64), float("nan"), device=dev)
    L.check(L.lib().f5b_op_attention_bwd(mode, q_d.data_ptr(), g_d.data_ptr(), L.ptr(l_d), batches, rows, heads,
                                         out.data_ptr(), 0, 0))
    return out.cpu()


def _attention_bwd_ref(qkv, d_out, kv_len, batches, rows, heads):
    """torch.autograd through the same softmax(q k^T / 8) v (what trainer.py:384 runs behind modules.py:499-508)."""
    x = qkv.double().clone().requires_grad_(True)
    y = x.view(batches, rows, 3, heads, 64)
    q, k, v = (y[:, :, i].transpose(1, 2) for i in range(3))
    s = q @ k.transpose(-1, -2) * 0.125
    if kv_len is not None:
        keep = torch.arange(rows)[None, :] < kv_len[:, None]
        s = s.masked_fill(~keep[:, None, None, :], float("-inf"))
    o = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(batches * rows, heads * 64)
    o.backward(d_out.double())
    return x.grad.float()


@pytest.mark.parametrize("mode,tol", [(0, 1e-4), (1, 2e-2)])
@pytest.mark.parametrize("batches,rows,heads,lens", [(1, 100, 2, None), (2, 333, 3, [333, 201]), (2, 940, 4, [940, 700]),
                                                      (1, 1024, 2, [128])])
def test_attention_backward(mode, tol, batches, rows, heads, lens):
    """dq | dk | dv of the attention core: fp32 CUDA-core kernels (mode 0) and the tcgen05 kernels (mode 1: bf16
    operands, P and dS rounded to bf16, fp32 accumulation) against autograd in float64.  Rows past an utterance's length
    carry no upstream gradient (the caller masks them, modules.py:540-542) and their keys take no part."""
    g = torch.Generator().manual_seed(rows + heads)
    qkv = torch.randn(batches * rows, 3 * heads * 64, generator=g)
    d_out = torch.randn(batches * rows, heads * 64, generator=g)
    kv = torch.tensor(lens) if lens else None
    if kv is not None:
        dead = (torch.arange(rows)[None, :] >= kv[:, None]).reshape(-1)
        d_out[dead] = 0.0
    if mode == 1:  # the kernel's operands are bf16: compare on what it was given
        qkv = qkv.bfloat16().float()
    got = _attention_bwd(mode, qkv, d_out, kv, batches, rows, heads)
    ref = _attention_bwd_ref(qkv, d_out, kv, batches, rows, heads)
    assert torch.isfinite(got).all()
    err = (got - ref).abs().max()
    assert err <= tol * max(1.0, float(ref.abs().max())), f"max err {err} (ref max {ref.abs().max()})"
    # relative Frobenius error per third: a wrong scale or a swapped block would pass a loose max-abs bound
    inner = heads * 64
    for i, name in enumerate(("dq", "dk", "dv")):
        a, b = got[:, i * inner:(i + 1) * inner], ref[:, i * inner:(i + 1) * inner]
        rel = float((a - b).norm() / b.norm())
        assert rel <= (1e-5 if mode == 0 else 1e-2), f"{name}: relative error {rel}"
