"""Print the SASS of one kernel of libewn_b200.so (substring match on the demangled name) and an opcode histogram.
usage: python tools/sass_fn.py k_emm_fused 'P5' [--dump]"""
import subprocess, sys, re, collections
so = "ewn_gym_b200/csrc/libewn_b200.so"
pats = [a for a in sys.argv[1:] if not a.startswith("--")]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", txt)
for b in blocks[1:]:
    name = b.split("\n", 1)[0].strip()
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    if all(p in dem for p in pats):
        ins = re.findall(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+)", b, re.M)
        print("==", dem, len(ins), "instructions")
        if "--dump" in sys.argv:
            for line in b.split("\n"):
                m = re.match(r"^\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
                if m: print(m.group(1), m.group(2).strip())
        else:
            print(collections.Counter(ins).most_common(30))
