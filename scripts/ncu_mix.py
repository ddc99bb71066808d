"""Instruction mix + hottest SASS lines of the FIRST kernel section of an `ncu --page source --csv --print-source sass` dump.
    python scripts/ncu_mix.py source.csv [top]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
ends = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
if len(ends) > 1:
    rows = rows[:ends[1]]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
print(rows[0][1][:100])
hdr = rows[1]
i_src = hdr.index('Source'); i_ex = hdr.index('Instructions Executed'); i_s = hdr.index('Warp Stall Sampling (All Samples)')
stall_cols = [(h, i) for i, h in enumerate(hdr) if h.startswith('stall_') and not h.endswith('(Not Issued)')]
agg = collections.Counter(); st = collections.Counter(); data = []
for r in rows[2:]:
    if len(r) <= i_ex: continue
    src = r[i_src].strip(); toks = src.split()
    op = toks[1] if toks and toks[0].startswith('@') and len(toks) > 1 else (toks[0] if toks else '')
    p = op.split('.')
    op = p[0] + ('.' + p[1] if p[0] in ('MUFU', 'LDTM', 'STS', 'F2FP', 'SYNCS', 'LDS', 'LDG', 'STG') and len(p) > 1 else '')
    agg[op] += int(r[i_ex] or 0); st[op] += int(r[i_s] or 0)
    data.append((int(r[i_s] or 0), int(r[i_ex] or 0), src, r))
tot = sum(agg.values()); ts = sum(st.values())
print("total warp-inst", tot, "samples", ts)
for k, v in agg.most_common(top):
    print(f"{k:22s} {v:10d} {100*v/tot:5.1f}%   stall samples {st[k]:6d} {100*st[k]/ts:5.1f}%")
for s, ex, src, r in sorted(data, key=lambda d: -d[0])[:top]:
    t2 = sorted(((int(r[i] or 0), h) for h, i in stall_cols), reverse=True)[:2]
    print(f"{s:6d} {100 * s / ts:5.1f}% ex={ex:8d} {src[:70]:70s} {t2}")
print(sorted({h: sum(int(r[i] or 0) for _, _, _, r in data) for h, i in stall_cols}.items(), key=lambda x: -x[1])[:8])
