"""First-contact diagnostics for the tcgen05 kernels on a real B200 (run under gpurun).  Each probe runs in its own
subprocess (a trap poisons the CUDA context) and appends a JSON line to gpurun_out/probe.jsonl."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "korean-f5-tts_b200"))
OUT = os.path.join(ROOT, "gpurun_out")


def emit(rec):
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, "probe.jsonl"), "a") as f:
        f.write(json.dumps(rec) + "\n")
    print(json.dumps(rec))


def probe_gemm(M, N, K, bn, pattern):
    import torch
    from f5b200 import _lib as L

    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(1)
    if pattern == "eye":
        a = torch.zeros(M, K)
        for i in range(min(M, K)):
            a[i, i] = 1.0
        w = (torch.arange(N * K).view(N, K) % 61).float() / 8
    else:
        a = torch.randn(M, K, generator=g)
        w = torch.randn(N, K, generator=g) * K ** -0.5
    out = torch.full((M, N), float("nan"), device=dev)
    a_d, w_d = a.to(dev), w.to(dev)
    rc = L.lib().f5b_op_gemm(1, a_d.data_ptr(), w_d.data_ptr(), None, M, N, K, 0, out.data_ptr(), bn, 0, 0)
    rec = dict(probe="gemm", M=M, N=N, K=K, bn=bn, pattern=pattern, rc=rc)
    if rc != 0:
        rec["err"] = L.lib().f5b_last_error().decode()
        return emit(rec)
    o = out.cpu()
    ref = (a.bfloat16().double() @ w.bfloat16().double().T).float()
    e = (o - ref).abs()
    e[torch.isnan(e)] = 1e9
    rec.update(max_err=float(e.max()), ref_max=float(ref.abs().max()), bad_frac=float((e > 1e-2 * (1 + ref.abs())).float().mean()),
               bad_rows=[int(i) for i in (e.max(1).values > 1e-2 * (1 + ref.abs().max())).nonzero().flatten()[:16]],
               bad_cols=[int(i) for i in (e.max(0).values > 1e-2 * (1 + ref.abs().max())).nonzero().flatten()[:16]],
               out00=[float(x) for x in o[0, :8]], ref00=[float(x) for x in ref[0, :8]],
               out10=[float(x) for x in o[1, :8]], ref10=[float(x) for x in ref[1, :8]])
    emit(rec)


def probe_attn(batches, rows, heads):
    import torch
    from f5b200 import _lib as L

    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(2)
    inner = heads * 64
    qkv = torch.randn(batches * rows, 3 * inner, generator=g)
    out = torch.full((batches * rows, inner), float("nan"), device=dev)
    q_d = qkv.to(dev)
    rc = L.lib().f5b_op_attention(1, q_d.data_ptr(), None, batches, rows, heads, out.data_ptr(), 0, 0)
    rec = dict(probe="attn", batches=batches, rows=rows, heads=heads, rc=rc)
    if rc != 0:
        rec["err"] = L.lib().f5b_last_error().decode()
        return emit(rec)
    x = qkv.bfloat16().double().view(batches, rows, 3, heads, 64)
    q, k, v = (x[:, :, i].transpose(1, 2) for i in range(3))
    p = torch.softmax(q @ k.transpose(-1, -2) * 0.125, -1)
    ref = (p @ v).transpose(1, 2).reshape(batches * rows, inner).float()
    # variant: what the output would be if V's swizzle / major-ness were ignored is not computed; report structure
    o = out.cpu()
    e = (o - ref).abs()
    e[torch.isnan(e)] = 1e9
    rec.update(max_err=float(e.max()), mean_err=float(e.mean()), bad_frac=float((e > 3e-2).float().mean()),
               bad_rows=[int(i) for i in (e.max(1).values > 3e-2).nonzero().flatten()[:16]],
               bad_cols=[int(i) for i in (e.max(0).values > 3e-2).nonzero().flatten()[:16]],
               out0=[float(t) for t in o[0, :8]], ref0=[float(t) for t in ref[0, :8]])
    emit(rec)


PROBES = [
    ("gemm", 128, 128, 64, 128, "eye"), ("gemm", 128, 128, 64, 128, "rand"), ("gemm", 128, 64, 128, 64, "rand"),
    ("gemm", 128, 256, 128, 256, "rand"), ("gemm", 300, 384, 512, 128, "rand"), ("gemm", 1880, 3072, 1024, 0, "rand"),
    ("attn", 1, 128, 1), ("attn", 1, 256, 1), ("attn", 1, 100, 2), ("attn", 2, 940, 16),
]

if __name__ == "__main__":
    if len(sys.argv) > 1:
        spec = json.loads(sys.argv[1])
        try:
            if spec[0] == "gemm":
                probe_gemm(*spec[1:])
            else:
                probe_attn(*spec[1:])
        except Exception as ex:  # noqa: BLE001
            emit(dict(probe=spec, exception=repr(ex)))
        sys.exit(0)
    for spec in PROBES:
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), json.dumps(spec)], timeout=180,
                               capture_output=True, text=True)
            if r.returncode != 0 or "probe" not in r.stdout:
                emit(dict(probe=spec, returncode=r.returncode, stdout=r.stdout[-1500:], stderr=r.stderr[-1500:]))
        except subprocess.TimeoutExpired:
            emit(dict(probe=spec, timeout=True))
