"""Hot SASS lines + stall reasons of one kernel from `ncu -i X.ncu-rep --page source --csv --print-source sass --kernel-id ...`.
    python scripts/ncu_hot_lines.py source.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
ends = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
if len(ends) > 1:
    rows = rows[:ends[1]]          # first kernel of the file only
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
print(rows[0][1][:120])
hdr = rows[1]
i_src = hdr.index('Source'); i_s = hdr.index('Warp Stall Sampling (All Samples)'); i_ex = hdr.index('Instructions Executed')
stall_cols = [(h, i) for i, h in enumerate(hdr) if h.startswith('stall_') and not h.endswith('_not_issued')]
data = [(int(r[i_s] or 0), int(r[i_ex] or 0), r[i_src].strip(), r) for r in rows[2:] if len(r) > i_ex]
tot = sum(d[0] for d in data); totex = sum(d[1] for d in data)
print("total samples", tot, "total warp-inst", totex)
for s, ex, src, r in sorted(data, key=lambda d: -d[0])[:top_n]:
    top = sorted(((int(r[i] or 0), h) for h, i in stall_cols), reverse=True)[:2]
    print(f"{s:6d} {100 * s / tot:5.1f}% ex={ex:8d} {src[:72]:72s} {top}")
agg = {h: sum(int(r[i] or 0) for _, _, _, r in data) for h, i in stall_cols}
print(sorted(agg.items(), key=lambda x: -x[1])[:10])
