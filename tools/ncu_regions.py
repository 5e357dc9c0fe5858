import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
warps = float(sys.argv[2]) if len(sys.argv) > 2 else 32768.0
cur=None; out={}
for r in rows:
    if not r: continue
    if r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if r[0] in ('Function Name','Line No'): continue
    if r[0].isdigit():
        try: ie=int(r[7])
        except: continue
        key=(cur,int(r[0]))
        if key in out: continue
        out[key]=ie
groups=[('core: bit intrinsics',('ewn_core.cuh',38,116)),('core: codec P5 (dead/dirs/x_of)',('ewn_core.cuh',135,150)),('core: unpack/pack',('ewn_core.cuh',173,216)),
('core: match/capture/win',('ewn_core.cuh',254,276)),('core: movers/cube_to_move',('ewn_core.cuh',280,301)),('core: legal/pick',('ewn_core.cuh',303,333)),
('core: apply/win_after/ply_action',('ewn_core.cuh',336,366)),('core: evaluate',('ewn_core.cuh',371,419)),('core: philox',('ewn_core.cuh',424,472)),('core: random_ply',('ewn_core.cuh',477,486)),
('core: step_one',('ewn_core.cuh',531,591)),('k: stage_rows',('ewn_kernels.cu',55,82)),('k: write_rows_dense',('ewn_kernels.cu',84,103)),('k: write_rows_sparse',('ewn_kernels.cu',105,116)),
('k: write_rows_compact',('ewn_kernels.cu',120,148)),('k: k_step prologue (288-311)',('ewn_kernels.cu',288,311)),('k: k_step action/args (312-327)',('ewn_kernels.cu',312,327)),('k: k_step scalar outputs (328-342)',('ewn_kernels.cu',328,342)),
('k: k_step stats (343-357)',('ewn_kernels.cu',343,357)),('k: k_step obs+term (358-385)',('ewn_kernels.cu',358,385)),('k: k_step epilogue',('ewn_kernels.cu',386,390))]
tot=sum(out.values()); acc=0
for name,(f,a,b) in groups:
    s=sum(v for (ff,l),v in out.items() if ff==f and a<=l<=b); acc+=s
    print(f"{s/warps:7.1f} {100*s/tot:5.1f}%  {name}")
print(f"{(tot-acc)/warps:7.1f} other;  total {tot/warps:.1f}")
for (ff,l),v in sorted(out.items()):
    if not any(ff==f and a<=l<=b for _,(f,a,b) in groups) and v/warps>1: print('   other:',ff,l,v/warps)
