"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 45
warps = float(sys.argv[3]) if len(sys.argv) > 3 else 32768.0
cur = None; out = {}
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] in ('Function Name', 'Line No'): continue
    if r[0].isdigit():
        try: ie = int(r[7]); smp = int(r[6])
        except Exception: continue
        key = (cur, int(r[0]))
        if key in out: continue
        out[key] = (ie, r[1][:100], smp, r[10])
tot = sum(v[0] for v in out.values())
print('total warp insts', tot, 'per warp', tot / warps)
for (f, l), v in sorted(out.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{v[0] / warps:6.1f} {100 * v[0] / tot:5.1f}% smp={v[2]:5d} thr={v[3]:>5} {f}:{l} {v[1]}")
