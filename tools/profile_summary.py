"""Turns one gpu_round's ncu outputs (gpurun_out/<tag>_*) into the committed summaries under profiles/.
    python tools/profile_summary.py r02a r02
"""
import csv, json, os, subprocess, sys

tag, name = sys.argv[1], sys.argv[2]
G, P = "gpurun_out", "profiles"
os.makedirs(P, exist_ok=True)

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum"]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    res = []
    for vals in rows[2:]:
        d = {h: (v, u) for h, v, u in zip(hdr, vals, units)}
        res.append(d)
    return res


lines = [f"# ncu summary {name} (gpurun tag {tag})", "",
         "Captured with `tools/gpu_round.sh` on one B200: `ncu --set full --clock-control none --import-source on` of",
         "`python bench.py --steps 2 --warmup 3 --no-cpu-baseline` (after the same command exited 0 without ncu).",
         "Per-launch numbers under ncu are cold-cache and serialised; bench.py's CUDA-event times are the reported ones.", ""]
traffic = {}
kernels = {}   # short kernel name -> hardware figures of one launch (bench.py: roofline.issue_active_pct ...)


def num(d, k):
    v, u = d.get(k, ("0", ""))
    try:
        return float(v) * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}.get(u, 1)
    except ValueError:
        return None


for rep, label in ((f"{G}/{tag}_render.ncu-rep", "render path"), (f"{G}/{tag}_tonemap.ncu-rep", "tonemap kernel")):
    if not os.path.exists(rep):
        continue
    for d in raw(rep):
        kn = d.get("Kernel Name", ("?", ""))[0]
        lines += [f"## {label}: `{kn}`", "", "| metric | value | unit |", "|---|---|---|"]
        for k in KEYS:
            if k in d:
                lines.append(f"| {k} | {d[k][0]} | {d[k][1]} |")
        stalls = sorted(((float(v[0]), h) for h, v in d.items() if h.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in h and v[0]), reverse=True)[:8]
        tot = sum(s for s, _ in stalls) or 1
        lines += ["", "Top warp-stall reasons (PC samples): " + ", ".join(f"{h.replace('smsp__pcsamp_warps_issue_stalled_', '')} {100 * s / tot:.0f}%" for s, h in stalls), ""]
        def mb(x):
            v, u = d.get(x, ("0", "byte"))
            v = float(v)
            return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        short = kn.split("(")[0].replace("void ", "").replace("rte::", "").split("<")[0]
        if short not in kernels or (num(d, "gpu__time_duration.sum") or 0) > (kernels[short]["time_us"] or 0):  # (render32: the whole-frame launch, not a band of rte_render_host)
            kernels[short] = {"kernel": kn.split("(")[0], "time_us": num(d, "gpu__time_duration.sum"), "registers": num(d, "launch__registers_per_thread"),
                              "grid": num(d, "launch__grid_size"), "block": num(d, "launch__block_size"), "inst_executed": num(d, "smsp__inst_executed.sum"),
                              "issue_active_pct": num(d, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                              "warps_active_pct": num(d, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                              "pipe_fma_pct": num(d, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                              "pipe_alu_pct": num(d, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                              "pipe_fp64_pct": num(d, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
                              "pipe_lsu_pct": num(d, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
                              "pipe_xu_pct": num(d, "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
                              "dram_bytes": (num(d, "dram__bytes_read.sum") or 0) + (num(d, "dram__bytes_write.sum") or 0),
                              "dram_pct": num(d, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")}
        traffic[kn.split("(")[0]] = {"dram_bytes_read": mb("dram__bytes_read.sum"), "dram_bytes_write": mb("dram__bytes_write.sum"),
                                     "time_ms_under_ncu": d["gpu__time_duration.sum"]}

ll = f"{G}/{tag}_launches.csv"
if os.path.exists(ll):
    rows = [r for r in csv.reader(open(ll)) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = {}
    for r in rows:
        if r is hdr or len(r) <= iv:
            continue
        try:
            t = float(r[iv].replace(",", ""))
        except ValueError:
            continue
        k = r[ik].split("(")[0]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1; a[1] += t
    tot = sum(a[1] for a in agg.values()) or 1
    lines += ["## launch list (`ncu --metrics gpu__time_duration.sum`, whole bench.py run)", "", "| kernel | launches | total ns | share |", "|---|---|---|---|"]
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        lines.append(f"| `{k}` | {n} | {t:.0f} | {100 * t / tot:.1f}% |")
    lines += ["", "(`at::` kernels are torch's L2-flush memsets / clones of the harness, not part of the step; `ffma*_peak_kernel` is the",
              "FP32-peak micro-benchmark; the step itself is `grid_build_kernel<0>` + `grid_scan_kernel` + `grid_build_kernel<1>` (the direction grids, rebuilt every step) + `render32_kernel`; `render_fast_kernel` / `tonemap64_kernel` are the exact path of the frame check.)", ""]
    import shutil
    shutil.copy(ll, f"{P}/{name}_launches.csv")
for f in ("bench.json", "bench_ref.json", "quick.txt", "gpu.txt", "pytest_gpu.txt", "smoke.txt", "c1.json", "tm_probe.txt", "grid_probe.txt"):
    src = f"{G}/{tag}_{f}"
    if os.path.exists(src):
        import shutil
        shutil.copy(src, f"{P}/{name}_{f}")
json.dump(traffic, open(f"{P}/{name}_traffic.json", "w"), indent=1)
json.dump(kernels, open(f"{P}/{name}_kernels.json", "w"), indent=1)
# SASS mnemonic counts per kernel of the shipped library (what proves sm_100a-native code paths; LDL / STL = local memory)
import collections, re
sass = subprocess.run(["cuobjdump", "-sass", "raytracingengine_b200/librte_b200.so"], capture_output=True, text=True).stdout
cur, counts = None, collections.OrderedDict()
for ln in sass.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
    if m and cur:
        op = m.group(1)
        counts[cur]["total"] += 1
        for key in ("FFMA2", "FFMA", "FMUL2", "UBLKCP", "SYNCS", "REDUX", "LDL", "STL", "DFMA", "DADD", "DMUL", "MUFU", "F2F", "LDS", "STS", "LDG", "STG", "ATOM", "RED", "SHFL", "VOTE", "BAR"):
            if op == key or (key in ("ATOM", "RED") and op.startswith(key)):
                counts[cur][key] += 1
with open(f"{P}/{name}_sass_counts.txt", "w") as f:
    f.write("SASS mnemonic counts per kernel / device function of raytracingengine_b200/librte_b200.so (cuobjdump -sass)\n")
    for k, c in counts.items():
        if c["total"] >= 40:
            f.write(f"{k}\n    " + "  ".join(f"{m}={c[m]}" for m in ("total", "FFMA", "FFMA2", "FMUL2", "DFMA", "DADD", "DMUL", "MUFU", "F2F", "LDS", "STS", "LDG", "STG", "LDL", "STL", "UBLKCP", "SYNCS", "REDUX", "SHFL", "VOTE", "ATOM", "RED", "BAR") if c[m]) + "\n")
open(f"{P}/{name}_ncu_summary.md", "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:80]))
