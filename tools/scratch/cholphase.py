import ctypes as C, numpy as np, katib_b200 as kb
from katib_b200 import _lib
L=_lib.lib()
ctx=kb.Context(0)
rng=np.random.Generator(np.random.PCG64(1))
n=512
M=rng.normal(size=(n,n)); A=M@M.T/n+np.diag(np.logspace(-3,2,n))
ctx.eigh(A)
out=(C.c_longlong*8)()
L.kb_dbg_chol_phases(out,1)
ctx.eigh(A)
L.kb_dbg_chol_phases(out,1)
names=["load","diag","trsm","store","sync1","update","sync2"]
tot=sum(out[:7])
print("cta0 cycles",tot,"= %.3f ms"%(tot/1.965e6))
for nm,v in zip(names,out): print("  %-8s %7.1f us  %4.1f%%"%(nm,v/1.965e3,100*v/tot))
