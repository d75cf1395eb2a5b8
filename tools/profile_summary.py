"""Turns one gpu_round's ncu outputs (gpurun_out/<tag>_*) into the committed summaries under profiles/.
    python tools/profile_summary.py r01p r01
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
for rep, label in ((f"{G}/{tag}_render.ncu-rep", "render kernel"), (f"{G}/{tag}_tonemap.ncu-rep", "tonemap kernel")):
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
              "FP32-peak micro-benchmark; the step itself is the primary-ray launch `render_fast_kernel<..., 1>`, the shading launch `<..., 2>` and `tonemap64_kernel`.)", ""]
    import shutil
    shutil.copy(ll, f"{P}/{name}_launches.csv")
for f in ("bench.json", "bench_ref.json", "quick.txt", "gpu.txt", "pytest_gpu.txt"):
    src = f"{G}/{tag}_{f}"
    if os.path.exists(src):
        import shutil
        shutil.copy(src, f"{P}/{name}_{f}")
json.dump(traffic, open(f"{P}/{name}_traffic.json", "w"), indent=1)
open(f"{P}/{name}_ncu_summary.md", "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:80]))
