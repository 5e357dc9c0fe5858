"""Control-flow skeleton of a kernel from an `ncu --import-source on` report: per-warp executed counts of every SASS
instruction, printed at branches / barriers / global memory instructions with the cost of the segment before them.
    python tools/ncu_sass_segments.py report.ncu-rep [warps_per_launch] [--full out.txt]"""
import csv, re, subprocess, sys
rep = sys.argv[1]
W = float(sys.argv[2]) if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else 32768.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = [r for r in csv.reader(out.splitlines()) if len(r) > 10 and r[0].startswith("0x")]
tot = sum(int(r[5]) for r in rows)
print(f"{len(rows)} SASS instructions, {tot / W:.1f} executed per warp")
if "--full" in sys.argv:
    with open(sys.argv[sys.argv.index("--full") + 1], "w") as f:
        for i, r in enumerate(rows):
            f.write(f"{i:4d} {int(r[5]) / W:5.2f} {float(r[8]):5.1f} {r[1]}\n")
cum = seg = 0.0
for i, r in enumerate(rows):
    c = int(r[5]) / W; cum += c; seg += c
    if re.search(r"BAR|BSYNC|BSSY|WARPSYNC|EXIT|VOTE|ATOM|RED|STG|LDG|BRA|DEPBAR", r[1]) and c > 0.01:
        print(f"{i:4d} {c:5.2f} {float(r[8]):5.1f} {r[1][:64]:64s} | seg={seg:6.1f} cum={cum:6.1f}")
        seg = 0.0
