"""Offline experiment (CPU, numpy/scipy): does a spatial order of the atoms INSIDE a link cell lower the number of 128-byte
lines a position-gather request of the once-per-pair list kernels touches?  (lever (b).)
A thermal FCC copper crystal (24^3 unit cells, sigma 0.08 A), link cells of rc + skin (Cu01), 4x4x4-cell work blocks, rows as
k_nl_build stores them (in-block pairs on the lower row only, cutoff part first, each part ascending); a request = one round
of 8 atoms x 4 lanes (G4) or 4 atoms x 8 lanes (G8); a line = 4 consecutive atoms of the sorted array (rq is 32 B per atom).
Orders compared: by gid (what the device uses), z-y-x sub-cells, Morton sub-cells (2 and 3 bits per axis), by x coordinate.
Result (profiles/r2_exp_incell_order.txt): 12.5-13.4 lines per request of ~25 lanes in every order -- a row's ~50 entries are
spread over 27 cells, 1.8 per cell, so consecutive entries sit in different cells and the order inside a cell is irrelevant."""
import numpy as np
from scipy.spatial import cKDTree
rng=np.random.default_rng(1)
n=24; lat=3.615; L=n*lat
rc=5.50679; rl=rc*1.1
# fcc
b=np.array([[.25,.25,.25],[.25,.75,.75],[.75,.25,.75],[.75,.75,.25]])
ix,iy,iz=np.meshgrid(np.arange(n),np.arange(n),np.arange(n),indexing='ij')
cells=np.stack([ix,iy,iz],-1).reshape(-1,1,3)
r=((cells+b[None])*lat).reshape(-1,3)
gid=np.arange(len(r))  # gid = basis + 4*(iz + nz*(iy+ny*ix)) in CoMD order
r=r+rng.normal(0,0.08,r.shape)
r%=L
nc=int(L/rl); cs=L/nc
c3=np.minimum((r/cs).astype(int),nc-1)
cell=c3[:,0]+nc*(c3[:,1]+nc*c3[:,2])
def morton(q,bits):
    k=np.zeros(len(q),np.int64)
    for b_ in range(bits):
        for a in range(3):
            k|=((q[:,a]>>b_)&1)<<(3*b_+a)
    return k
frac=(r/cs)-c3
def order(kind):
    if kind=='gid': key=gid
    elif kind=='morton2': key=morton((frac*4).astype(int),2)*len(r)+gid
    elif kind=='morton3': key=morton((frac*8).astype(int),3)*len(r)+gid
    elif kind=='zyx': key=(frac[:,2]*4).astype(int)*16+(frac[:,1]*4).astype(int)*4+(frac[:,0]*4).astype(int); key=key*len(r)+gid
    elif kind=='x': key=(frac[:,0]*1e6).astype(np.int64)
    return np.lexsort((key,cell))
tree=cKDTree(r,boxsize=L)
pairs=tree.query_pairs(rl,output_type='ndarray')
bx=4
blk3=c3//bx
blk=blk3[:,0]+100*(blk3[:,1]+100*blk3[:,2])
for kind in ['gid','zyx','morton2','morton3','x']:
    o=order(kind); rank=np.empty(len(r),int); rank[o]=np.arange(len(r))
    i=rank[pairs[:,0]]; j=rank[pairs[:,1]]
    same=blk[pairs[:,0]]==blk[pairs[:,1]]
    lo=np.minimum(i,j); hi=np.maximum(i,j)
    # rows: in-block pair stored on lower only; cross-block both
    ri=np.concatenate([lo,hi[~same]]); rj=np.concatenate([hi,lo[~same]])
    d=np.linalg.norm((lambda v:(v-L*np.round(v/L)))(r[o][ri]-r[o][rj]),axis=1)
    near=d<=rc
    # order within row: near first then far, each ascending index
    s=np.lexsort((rj,~near,ri))
    ri=ri[s]; rj=rj[s]
    # position in row
    start=np.searchsorted(ri,np.arange(len(r)))
    pos=np.arange(len(ri))-start[ri]
    # warp request = 8 consecutive atoms (ri//8) x round (pos//4): count distinct lines rj//4
    req=(ri//8)*1000+(pos//4)
    key=req*10**7+(rj//4)
    nreq=len(np.unique(req)); nlines=len(np.unique(key))
    # 8-lane G: 4 atoms x 8 entries
    req8=(ri//4)*1000+(pos//8); key8=req8*10**7+(rj//4)
    print(kind,'entries/atom %.1f'%(len(ri)/len(r)),'G4: lines/request %.2f lanes/request %.1f'%(nlines/nreq,len(ri)/nreq),' G8: %.2f %.1f'%(len(np.unique(key8))/len(np.unique(req8)),len(ri)/len(np.unique(req8))))
