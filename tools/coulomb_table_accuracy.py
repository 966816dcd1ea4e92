import numpy as np, mpmath as mp
from scipy.special import erfc
mp.mp.dps=40
alpha=0.162157
def build(D,M,alpha,e_lo,e_hi):
    NS=1<<M; ld=np.longdouble
    nodes=np.cos(np.pi*(2*np.arange(D+1)+1)/(2.0*(D+1))).astype(ld)
    tabs={}
    for e in range(e_lo,e_hi):
        for j in range(NS):
            s0=ld(2.0)**e*(1+ld(j)/NS); w=ld(2.0)**e/NS
            sv=s0+(nodes+1)*ld(0.5)*w
            # exact values in mp
            f=np.array([ld(str(mp.erfc(mp.mpf(alpha)*mp.sqrt(mp.mpf(str(x))))/mp.sqrt(mp.mpf(str(x))))) for x in sv])
            V=np.vander(nodes,D+1,increasing=True).astype(ld)
            c=np.linalg.solve(V.astype(np.float64),f.astype(np.float64)) # rough; refine in ld
            # iterative refinement in long double
            for _ in range(3):
                r=f-V@c.astype(ld); c=c.astype(ld)+np.linalg.solve(V.astype(np.float64),r.astype(np.float64)).astype(ld)
            tabs[(e,j)]=np.array(c,dtype=np.float64)
    return tabs
def evalt(tabs,D,M,s):
    NS=1<<M
    m,e=np.frexp(s); e=e-1; m=m*2  # s = m*2^e, m in [1,2)
    j=np.floor((m-1)*NS).astype(int); j=np.minimum(j,NS-1)
    t=2*((m-1)*NS-j)-1
    out=np.zeros_like(s)
    for i in range(len(s)):
        c=tabs[(int(e[i]),int(j[i]))]; p=c[D]
        for d in range(D-1,-1,-1): p=p*t[i]+c[d]
        out[i]=p
    return out
rng=np.random.default_rng(0)
r=rng.uniform(1.5,37,4000); s=r*r
exact=np.array([float(mp.erfc(mp.mpf(alpha)*mp.mpf(float(x)))/mp.mpf(float(x))) for x in r])
# recompute exact from s to be consistent
exact=np.array([float(mp.erfc(mp.mpf(alpha)*mp.sqrt(mp.mpf(float(x))))/mp.sqrt(mp.mpf(float(x)))) for x in s])
for D,M in ((9,3),(9,4),(10,3),(11,3),(8,4),(7,4)):
    tabs=build(D,M,alpha,1,11)
    v=evalt(tabs,D,M,s)
    err=np.abs(v-exact); rel=err/exact
    print(D,M,"abs max %.2e rms %.2e | rel max(r<10) %.2e rel rms(r<10) %.2e | glibc rel rms %.2e"%(err.max(),np.sqrt((err**2).mean()),rel[r<10].max(),np.sqrt((rel[r<10]**2).mean()), np.sqrt((((erfc(alpha*np.sqrt(s))/np.sqrt(s)-exact)/exact)[r<10]**2).mean())))
