"""In-situ timeline of the graph-replayed sampling loop (the view a system profiler would give).

Needs the trace build of the library:  F5B_KTRACE=1 python korean-f5-tts_b200/build.py --force
Every CTA of the per-step kernels records globaltimer at entry, after its programmatic-dependency wait and at exit.
Because a dependent grid passes its wait only after the previous grid has completed, launches are recovered by sorting
the records by the post-wait stamp and cutting where a stamp exceeds every exit seen so far.

Per launch:   gap   = first post-wait stamp - last exit of the previous launch   (launch / dependency latency)
              ramp  = last post-wait stamp - first post-wait stamp                (CTAs becoming resident)
              busy  = last exit - first post-wait stamp                           (the launch's own span)
              tail  = last exit - median exit                                     (imbalance at the end)
Prints one evaluation (between two euler_xproj launches) kernel by kernel and the per-kernel totals.
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "korean-f5-tts_b200"))
import f5b200  # noqa: E402
from f5b200 import _lib as L  # noqa: E402

NAMES = {1: "tc_gemm", 2: "tc_attn", 4: "ln_mod", 8: "euler_xproj"}


def main():
    dev = torch.device("cuda", 0)
    dit, _ = bench.synth_state(0)
    cfm = f5b200.CFM(transformer=dit.to(dev), mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()
    kw = dict(duration=bench.N_TOTAL, steps=bench.NFE, cfg_strength=bench.CFG, sway_sampling_coef=bench.SWAY,
              return_trajectory=False)
    ins = [tuple(t.to(dev) for t in bench.synth_inputs(i)) for i in range(3)]
    for cond, text, y0 in ins[:2]:
        cfm.sample(cond=cond, text=text, y0=y0, **kw)
    torch.cuda.synchronize()
    cap = 1 << 20
    n = C.c_int64(cap)
    L.check(L.lib().f5b_debug_ktrace(0, cap, None, None, 0))
    cond, text, y0 = ins[2]
    cfm.sample(cond=cond, text=text, y0=y0, **kw)
    torch.cuda.synchronize()
    out = np.zeros((cap, 8), dtype=np.int64)
    L.check(L.lib().f5b_debug_ktrace(1, cap, out.ctypes.data, C.addressof(n), 0))
    rec = out[: min(n.value, cap)]
    print(f"{n.value} CTA records")
    rec = rec[np.argsort(rec[:, 1], kind="stable")]
    launches = []  # [kid, ctas, first_entry, first_wait, last_wait, median_exit, last_exit]
    cur = []
    max_exit = 0
    for r in rec:
        if cur and r[1] >= max_exit:
            launches.append(cur)
            cur = []
        cur.append(r)
        max_exit = max(max_exit, int(r[2]))
    launches.append(cur)
    rows = []
    prev_end = None
    for ls in launches:
        a = np.array(ls)
        kid = int(a[0, 3] >> 32)
        fw, lw, le = int(a[:, 1].min()), int(a[:, 1].max()), int(a[:, 2].max())
        rows.append(dict(kid=kid, ctas=len(ls), entry_lead=(fw - int(a[:, 0].min())) / 1e3,
                         gap=(fw - prev_end) / 1e3 if prev_end else 0.0, ramp=(lw - fw) / 1e3, busy=(le - fw) / 1e3,
                         tail=(le - float(np.median(a[:, 2]))) / 1e3, t0=fw, t1=le,
                         # GEMM phases, median over CTAs: operands landed, MMAs issued, accumulator complete, stores issued
                         ph=[float(np.median(a[:, 4] - a[:, 1])) / 1e3, float(np.median(a[:, 5] - a[:, 4])) / 1e3,
                             float(np.median(a[:, 6] - a[:, 5])) / 1e3, float(np.median(a[:, 7] - a[:, 6])) / 1e3,
                             float(np.median(a[:, 2] - a[:, 7])) / 1e3] if kid == 1 else
                         [float(np.median(a[:, 4] - a[:, 1])) / 1e3, float(np.median(a[:, 5] - a[:, 4])) / 1e3,
                          float(np.median(a[:, 2] - a[:, 5])) / 1e3] if kid == 8 else None))
        prev_end = le
    eul = [i for i, r in enumerate(rows) if r["kid"] == 8]
    print(f"{len(rows)} launches, {len(eul)} euler_xproj")
    a, b = eul[3], eul[4]
    span = (rows[b]["t0"] - rows[a]["t0"]) / 1e3
    print(f"\none evaluation (launches {a}..{b}): {span:.1f} us wall")
    print(f"{'kernel':12s} {'ctas':>5s} {'gap':>6s} {'ramp':>6s} {'busy':>7s} {'tail':>6s} {'pre-wait':>8s}")
    for r in rows[a : a + 16] + rows[b - 4 : b + 1]:
        ph = "" if r["ph"] is None else ("   fill %.2f  issue %.2f  drain %.2f  epilogue %.2f  store %.2f" % tuple(r["ph"])
                                          if len(r["ph"]) == 5 else "   stage %.2f  contract %.2f  add+store %.2f" % tuple(r["ph"]))
        print(f"{NAMES.get(r['kid'], r['kid']):12s} {r['ctas']:5d} {r['gap']:6.2f} {r['ramp']:6.2f} {r['busy']:7.2f} "
              f"{r['tail']:6.2f} {r['entry_lead']:8.2f}{ph}")
    agg = {}
    for r in rows[a:b]:
        k = (NAMES.get(r["kid"], r["kid"]), r["ctas"])
        x = agg.setdefault(k, [0, 0.0, 0.0, 0.0])
        x[0] += 1
        x[1] += r["busy"]
        x[2] += r["gap"]
        x[3] += r["tail"]
    print(f"\n{'kernel':12s} {'ctas':>5s} {'n':>4s} {'busy_us':>9s} {'avg':>7s} {'gap_us':>8s} {'avg':>6s} {'tail avg':>8s}")
    tb = tg = 0.0
    for k, (c, busy, gap, tail) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k[0]:12s} {k[1]:5d} {c:4d} {busy:9.1f} {busy / c:7.2f} {gap:8.1f} {gap / c:6.2f} {tail / c:8.2f}")
        tb += busy
        tg += gap
    print(f"busy {tb:.1f} us + gaps {tg:.1f} us = {tb + tg:.1f} us of {span:.1f} us")


if __name__ == "__main__":
    main()
