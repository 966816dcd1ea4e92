import sys, argparse, numpy as np
sys.path.insert(0,'/root/repo')
import bench
s=bench.make_system(argparse.Namespace(framework=50000,guests=200))
r_fw=int(np.flatnonzero(s.is_active!=1)[0]); mo,ao=s.mol_offsets()
n=50000
pos=(s.com[mo[r_fw]][None,:]+s.offsets[ao[r_fw]:ao[r_fw]+n])
L=s.box_len; lo=s.box_lo
u=((pos-lo)/L)%1.0
bits=1
while bits<10 and (1<<(3*bits))*2<n: bits+=1
nc=1<<bits
cell=np.minimum((u*nc).astype(np.int64),nc-1)
def spread(v):
    r=np.zeros_like(v)
    for b in range(10): r|=((v>>b)&1)<<(3*b)
    return r
code=spread(cell[:,0])|(spread(cell[:,1])<<1)|(spread(cell[:,2])<<2)
order=np.argsort(code,kind='stable')
w=lo+u*L
wt=w[order]
nt=(n+31)//32
ext=[];cen=[]
for t in range(nt):
    p=wt[t*32:(t+1)*32]; ext.append((p.max(0)-p.min(0))/2); cen.append((p.max(0)+p.min(0))/2)
ext=np.array(ext);cen=np.array(cen)
print("bits",bits,"cells/axis",nc,"cell size",L[0]/nc,"tiles",nt)
print("half-extent mean",ext.mean(0),"max",ext.max(0),"median",np.median(ext,0))
reach=6/s.alpha; R=reach+1.5
rng=np.random.default_rng(0)
cnts=[];atoms=[]
for _ in range(200):
    c=lo+rng.random(3)*L
    d=cen-c; d-=L*np.round(d/L); d=np.maximum(0,np.abs(d)-ext)
    hit=(d**2).sum(1)<=R*R; cnts.append(hit.sum())
    dd=w-c; dd-=L*np.round(dd/L); atoms.append(((dd**2).sum(1)<=R*R).sum())
print("reach",reach,"tiles hit mean",np.mean(cnts),"=",np.mean(cnts)*32,"lane-atoms vs atoms in sphere",np.mean(atoms), "efficiency",np.mean(atoms)/(np.mean(cnts)*32))

def kd_tiles(idx, pts, out, cap=32):
    if len(idx) <= cap:
        out.append(idx); return
    p = pts[idx]; ax = int(np.argmax(p.max(0) - p.min(0)))
    # split so that the left part is a multiple of cap (full tiles) close to the median
    n = len(idx); left = max(cap, int(round(n / 2 / cap)) * cap)
    if left >= n: left = n // 2
    o = np.argsort(p[:, ax], kind='stable')
    kd_tiles(idx[o[:left]], pts, out, cap); kd_tiles(idx[o[left:]], pts, out, cap)
tl=[]; kd_tiles(np.arange(n), w, tl)
ext2=np.array([(w[t].max(0)-w[t].min(0))/2 for t in tl]); cen2=np.array([(w[t].max(0)+w[t].min(0))/2 for t in tl])
sizes=np.array([len(t) for t in tl])
print("kd tiles",len(tl),"fill",sizes.mean()/32,"half-extent mean",ext2.mean(0),"max",ext2.max(0))
rng=np.random.default_rng(0); cnts2=[]
for _ in range(200):
    c=lo+rng.random(3)*L
    d=cen2-c; d-=L*np.round(d/L); d=np.maximum(0,np.abs(d)-ext2)
    cnts2.append(((d**2).sum(1)<=R*R).sum())
print("kd tiles hit mean",np.mean(cnts2),"vs morton",np.mean(cnts))
