import numpy as np, mpmath as mp
mp.mp.prec=120
lo,hi=-1/3,1/3
N=600
k=np.arange(N); x=np.cos(np.pi*(k+0.5)/N); r=(lo+hi)/2+(hi-lo)/2*x
T=np.array([float((mp.log1p(mp.mpf(v))-mp.mpf(v)+mp.mpf(v)**2/2)/mp.mpf(v)**3) for v in r])
W=np.array([float(mp.mpf(v)**3/mp.log1p(mp.mpf(v))) for v in r])   # relative-error weight
DEG=7
V=np.vander(r,DEG+1,increasing=True)
lw=np.ones(N)
for it in range(60):   # Lawson iterations toward minimax
    A=V*(W*np.sqrt(lw))[:,None]; b=T*W*np.sqrt(lw)
    c,_,_,_=np.linalg.lstsq(A,b,rcond=None)
    err=np.abs((V@c-T)*W)
    lw=lw*err; lw/=lw.sum()
print("max rel err (f64 coeffs)",err.max())
c32=c.astype(np.float32)
print([repr(float(v)) for v in c32])
# evaluate f32 algorithm
f=np.float32
def fma(a,b,c): return (a.astype(np.float64)*b.astype(np.float64)+c.astype(np.float64)).astype(f)
def mylog(x):
    i=x.view(np.int32)
    e=(i-np.int32(0x3f2aaaab))&np.int32(-8388608)   # 0xff800000
    m=(i-e).view(f)
    fe=e.astype(f)          # int->float (exact for multiples of 2^23 up to ±127*2^23? needs 8 bits mantissa: exact)
    rr=(m-f(1.0)).astype(f)
    s=(rr*rr).astype(f)
    q=np.full_like(rr,c32[DEG])
    for j in range(DEG-1,-1,-1): q=fma(q,rr,np.full_like(rr,c32[j]))
    p=fma(q,rr,np.full_like(rr,f(-0.5)))
    res=fma(p,s,rr)
    res=fma(fe,np.full_like(rr,f(0.693147182464599609375*2.0**-23)),res)
    return res
rng=np.random.RandomState(0)
worst=0
for ex in range(-40,41):
    m=rng.randint(0,1<<23,400000).astype(np.int32)
    xi=((ex+127)<<23)|m
    x=xi.astype(np.int32).view(f)
    got=mylog(x).astype(np.float64)
    ref=np.log(x.astype(np.float64))
    ulp=np.abs(np.spacing(ref.astype(f))).astype(np.float64)
    e=np.abs(got-ref)/ulp
    e[ref==0]=np.abs(got[ref==0])
    worst=max(worst,e.max())
print("max ulp err",worst)
# near 1
x=(f(1.0)+np.arange(-200000,200000).astype(f)*f(2**-24)).astype(f)
got=mylog(x).astype(np.float64); ref=np.log(x.astype(np.float64))
ok=ref!=0
print("near 1 max rel", np.max(np.abs(got[ok]-ref[ok])/np.abs(ref[ok])), "log(1)=",mylog(np.array([1.0],f)))
