"""Aggregates an `ncu --page source --csv` dump by code region: samples, stall reasons, executed instructions."""
import csv, sys
path = sys.argv[1]; B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
rows = list(csv.reader(open(path)))
hdr = rows[1]; data = rows[2:]
isrc = hdr.index('Source'); ins = hdr.index('# Samples'); iex = hdr.index('Instructions Executed'); ith = hdr.index('Avg. Threads Executed')
st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[ins] or 0) for r in data); totex = sum(int(r[iex] or 0) for r in data)
print('instrs', len(data), 'samples', tot, 'exec', totex)
print({h: sum(int(r[hdr.index(h)] or 0) for r in data) for h in st if sum(int(r[hdr.index(h)] or 0) for r in data) > tot * 0.02})
for b in range(0, len(data), B):
    chunk = data[b:b + B]
    s = sum(int(r[ins] or 0) for r in chunk); e = sum(int(r[iex] or 0) for r in chunk)
    if s > tot * 0.012:
        ops = {}
        for r in chunk:
            t = r[isrc].split()
            o = t[1] if t and t[0].startswith('@') and len(t) > 1 else (t[0] if t else '')
            ops[o] = ops.get(o, 0) + 1
        top = sorted(ops.items(), key=lambda x: -x[1])[:4]
        stl = {h[6:]: sum(int(r[hdr.index(h)] or 0) for r in chunk) for h in st}
        stl = sorted(stl.items(), key=lambda x: -x[1])[:3]
        thr = sum(float(r[ith] or 0) * int(r[iex] or 0) for r in chunk) / max(1, e)
        print(b, 'samples %.1f%%' % (100 * s / tot), 'exec %.1f%%' % (100 * e / totex), 'thr %.1f' % thr, stl, top)
