#!/usr/bin/env python
"""Windowed list schedules for the lj/cut gather (no GPU needed; development tool like line_model.py): the lanes of a 32-atom
tile sweep their common neighbourhood in the same order; a lane whose next neighbor lies more than W sweep units ahead of
the slowest lane idles for that step.  Counts steps per warp and distinct 128-byte lines per step on a thermalised melt.
Result: W = 48: 1353 L1 wavefronts per warp at 132 steps against 1695 at 80 steps unwindowed."""
import sys
sys.path.insert(0, str(__import__('pathlib').Path(__file__).resolve().parents[1]))
import numpy as np
from scipy.spatial import cKDTree
import oracle
cells=16
s=oracle.make_lj_system(cells); s.setup(); s.run(100)
x=s.atoms()["x"].copy()
L=cells*(4/0.8442)**(1/3); x%=L; n=len(x)
tree=cKDTree(x,boxsize=L)
pr=tree.query_pairs(2.8,output_type="ndarray")
I=np.concatenate([pr[:,0],pr[:,1]]); J=np.concatenate([pr[:,1],pr[:,0]])
nb=int(L/2.8); bs=L/nb
b=np.minimum((x/bs).astype(int),nb-1)
binkey=(b[:,2]*nb+b[:,1])*nb+b[:,0]
f=np.minimum(((x-b*bs)/(bs/4)).astype(int),3)
k=np.zeros(n,dtype=np.int64)
for bit in range(2):
    for d in range(3):
        k|=((f[:,d]>>bit)&1).astype(np.int64)<<(3*bit+d)
perm=np.lexsort((k,binkey)); rank=np.empty(n,dtype=np.int64); rank[perm]=np.arange(n)
ri,rj=rank[I],rank[J]
# order per atom: (dz,dy) rows, dxb, index  (as the kernel emits) -> use a monotone key for the sweep
dy=(b[J,1]-b[I,1]+nb+1)%nb; dz=(b[J,2]-b[I,2]+nb+1)%nb; dxb=(b[J,0]-b[I,0]+nb+1)%nb
# sweep coordinate: position of j in i's 27-bin stencil order = ((dz*3+dy)*3+dxb)*64 + index within bin
binstart=np.searchsorted(np.sort(binkey),np.arange(nb**3))
jin=rj-binstart[binkey[J]]
sweep=((dz*3+dy)*3+dxb)*64+jin           # per-atom-relative sweep position; lanes of a tile in different bins: comparable approx
order=np.lexsort((sweep,ri)); ri,rj,sweep=ri[order],rj[order],sweep[order]
start=np.searchsorted(ri,np.arange(n+1))
def simulate(W, ntiles=400, unit=4):
    steps=0; lines=0; act=0
    for t in range(ntiles):
        lanes=range(t*32,min(t*32+32,n))
        lists=[(rj[start[a]:start[a+1]], sweep[start[a]:start[a+1]]//unit) for a in lanes]
        p=[0]*len(lists)
        while True:
            rem=[i for i,(l,_) in enumerate(lists) if p[i]<len(l)]
            if not rem: break
            lo=min(lists[i][1][p[i]] for i in rem)
            ls=set()
            for i in rem:
                if lists[i][1][p[i]]<lo+W:
                    ls.add(lists[i][0][p[i]]//4); p[i]+=1; act+=1
            steps+=1; lines+=len(ls)
    return steps/ntiles, lines/steps, lines/ntiles, act/steps
for W in (10**9, 400, 200, 120, 80, 48, 32, 16):
    st,lps,lpw,a=simulate(W)
    print(f"W={W:>10}: steps/warp {st:6.1f} lines/step {lps:5.1f} lines/warp {lpw:7.0f} lanes/step {a:5.1f}  cycles: L1 {lpw+st:6.0f} FP64 {st*8.5:6.0f}")
