import time, numpy as np, katib_b200 as kb, sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import bench
kinds, lows, highs, steps, nch = bench.workload_space()
ctx = kb.Context(0)
ctx.space_set(kinds, lows, highs, steps, nch)
for n in (100, 1000, 10000, 100000):
    vals, loss = bench.workload_history(kinds, lows, highs, steps, nch, n=n)
    for merge in (0, 1):
        ctx.tpe_set_option(1, merge)
        ctx.tpe_set_history(vals, loss); ctx.tpe_get_parzen(0, 0)
        ts = []
        for _ in range(10):
            t0 = time.perf_counter(); ctx.tpe_set_history(vals, loss); ctx.tpe_get_parzen(0, 0); ts.append(time.perf_counter() - t0)
        ctx.profile(True); ctx.tpe_set_history(vals, loss); ctx.tpe_get_parzen(0, 0); ms, k = ctx.profile_read(5); ctx.profile(False)
        print(f"n={n:6d} merge={merge} set_history + sync: {1e3 * np.median(ts):8.3f} ms (50 parameters); build kernels {ms:.3f} ms")
