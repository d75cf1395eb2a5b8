"""Per-source-line instruction / stall-sample shares of one kernel from an ncu capture (developer tool).

    python tools/ncu_lines.py gpurun_out/x.ncu-rep render32 [top_n]

ncu's source page (CSV) has per-SASS-address counts; `nvdisasm -g` of the cubin inside librte_b200.so has the source line of
every SASS offset.  The two are joined by instruction order.  (The .so must be the build the capture was taken from.)
"""
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 45

out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
kernel = rows[0][1]
hdr = rows[1]
ia, ie, ism = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
sass = []
for r in rows[2:]:
    if len(r) != len(hdr) or r[0] == "Kernel Name":
        break
    sass.append((int(r[ie] or 0), int(r[ism] or 0), r[1].strip()))

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "raytracingengine_b200", "librte_b200.so")], cwd=tmp, capture_output=True)
lines = None
short = re.sub(r"\(.*", "", kernel).replace("void ", "").replace("rte::", "").split("<")[0]
for cub in glob.glob(os.path.join(tmp, "*.cubin")):
    txt = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
    # sections: .text.<mangled>
    parts = re.split(r"\n\s*\.section\s+\.text\.", txt)
    for part in parts[1:]:
        name = part.split(",")[0]
        if short in name:
            cur, lst = ("?", 0), []
            for ln in part.splitlines():
                m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
                if m:
                    cur = (os.path.basename(m.group(1)), int(m.group(2)))
                    continue
                if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
                    lst.append(cur)
            if len(lst) == len(sass):
                lines = lst
if lines is None:
    sys.exit(f"no cubin function with {len(sass)} instructions matches {kernel}")

tot_i = sum(s[0] for s in sass) or 1
tot_s = sum(s[1] for s in sass) or 1
agg = {}
for (n, smp, _), key in zip(sass, lines):
    a = agg.setdefault(key, [0, 0, 0])
    a[0] += n; a[1] += smp; a[2] += 1
src_cache = {}


def src(f, l):
    if f not in src_cache:
        p = os.path.join(ROOT, "raytracingengine_b200", "csrc", f)
        src_cache[f] = open(p).read().splitlines() if os.path.exists(p) else []
    s = src_cache[f]
    return s[l - 1].strip()[:100] if 0 < l <= len(s) else ""


print(f"{kernel}: {len(sass)} SASS instructions, {tot_i} executed, {tot_s} samples")
byfile = {}
for (f, l), (n, smp, k) in agg.items():
    b = byfile.setdefault(f, [0, 0])
    b[0] += n; b[1] += smp
for f, (n, smp) in sorted(byfile.items(), key=lambda x: -x[1][0]):
    print(f"  {f:22s} instr {100 * n / tot_i:5.1f}%  samples {100 * smp / tot_s:5.1f}%")
print("  inst%  samp%  sass  line")
for (f, l), (n, smp, k) in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    print(f"  {100 * n / tot_i:5.1f}  {100 * smp / tot_s:5.1f}  {k:4d}  {f}:{l}  {src(f, l)}")
