import time, numpy as np, katib_b200 as kb
ctx=kb.Context(0)
rng=np.random.Generator(np.random.PCG64(1))
for n in (64,128,256,512,1024):
    M=rng.normal(size=(n,n)); A=M@M.T/n+np.diag(np.logspace(-3,2,n))
    ctx.eigh(A)
    t0=time.perf_counter(); w,V=ctx.eigh(A); dt=time.perf_counter()-t0
    ctx.profile(True); ctx.eigh(A); ms,_=ctx.profile_read(9); ctx.profile(False)
    sw=ctx.eigh_sweeps()
    print(n, "wall %.1f ms kernel %.1f ms sweeps %d -> %.2f us/step"%(dt*1e3, ms, sw, ms*1e3/(sw*(n-1+(n&1)))))
# warm-start behaviour inside CMA-ES
n,lam=512,4096
ctx.cmaes_create(np.full(n,3.0),2.0,lam)
for g in range(6):
    X=ctx.cmaes_ask(lam, rng.normal(size=(lam,n))); sw=ctx.eigh_sweeps()
    f=np.sum((X*np.logspace(0,3,n))**2,axis=1); ctx.cmaes_tell(X,f)
    print("gen",g,"sweeps",sw)
