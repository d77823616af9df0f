"""Brief of an .ncu-rep: per captured launch the headline metrics; with --src N the hottest SASS lines of launch N."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ["Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max",
        "launch__registers_per_thread", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "lts__t_sectors_op_red.sum",
        "lts__t_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__inst_executed_pipe_xu.sum.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
idx = [hdr.index(w) for w in want if w in hdr]
for n, r in enumerate(rows[2:]):
    print(n, {hdr[i].split(".")[0][-28:] + ("." + hdr[i].split(".")[1] if "." in hdr[i] else ""): r[i] for i in idx})
if "--src" in sys.argv:
    k = int(sys.argv[sys.argv.index("--src") + 1])
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    kern, cur = [], None
    for r in csv.reader(io.StringIO(src)):
        if r and r[0] == "Kernel Name":
            cur = []; kern.append(cur); continue
        if cur is not None: cur.append(r)
    h, body = kern[k][0], kern[k][1:]
    isamp, isrc, iex = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
    sc = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
    tot = sum(int(r[isamp]) for r in body)
    agg = {}
    for r in body:
        for c in sc: agg[h[c]] = agg.get(h[c], 0) + int(r[c])
    print("samples", tot, "instr", len(body), sorted(agg.items(), key=lambda x: -x[1])[:8])
    for i in sorted(sorted(range(len(body)), key=lambda i: -int(body[i][isamp]))[:int(sys.argv[sys.argv.index("--src") + 2]) if len(sys.argv) > sys.argv.index("--src") + 2 else 25]):
        r = body[i]; st = {h[c][6:]: int(r[c]) for c in sc if int(r[c]) > 0}
        print(i, r[isamp], r[iex], r[isrc].strip()[:64], dict(sorted(st.items(), key=lambda x: -x[1])[:3]))
