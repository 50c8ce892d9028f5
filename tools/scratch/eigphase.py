import ctypes as C, numpy as np, katib_b200 as kb, os
from katib_b200 import _lib
L=_lib.lib()
ctx=kb.Context(0)
rng=np.random.Generator(np.random.PCG64(1))
n=512
M=rng.normal(size=(n,n)); A=M@M.T/n+np.diag(np.logspace(-3,2,n))
ctx.eigh(A)
out=(C.c_longlong*8)()
L.kb_dbg_eigh_phases(out,1)
ctx.profile(True); ctx.eigh(A); ms,_=ctx.profile_read(9); ctx.profile(False)
L.kb_dbg_eigh_phases(out,1)
sw=ctx.eigh_sweeps()
names=["wait","load","gram","small","apply","store","gridsync"]
tot=sum(out[:7])
print("M",os.environ.get("KB_EIGH_M"),"kernel ms",ms,"sweeps",sw,"cta0 cycles",tot, "= %.2f ms @1.965GHz"%(tot/1.965e6))
for nm,v in zip(names,out): print("  %-8s %6.2f ms  %4.1f%%"%(nm,v/1.965e6,100*v/tot))
