"""Per-source-line instruction / stall-sample shares from `ncu --page source --print-source cuda,sass --csv`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; agg = {}
for r in rows:
    if len(r) >= 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if len(r) < 9 or r[0] in ('Line No', 'Function Name'): continue
    if r[0] != '' and r[2] == '-':
        try: ln = int(r[0]); inst = float(r[7]); samples = float(r[6]); thr = float(r[8])
        except ValueError: continue
        agg[(cur, ln)] = (inst, samples, thr, r[1].strip()[:100])
tot = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values()); tt = sum(v[2] for v in agg.values())
print(f"total warp inst {tot:.4g}  thread inst {tt:.4g}  avg threads {tt / tot:.1f}  samples {ts:.0f}")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top_n]:
    print("%-18s %4d inst %5.2f%% smp %5.2f%% thr %4.1f | %s" % (f, l, 100 * v[0] / tot, 100 * v[1] / ts, v[2] / max(v[0], 1), v[3]))
