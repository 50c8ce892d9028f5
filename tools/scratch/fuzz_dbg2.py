import sys, os, numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import katib_b200 as kb, _oracle as O
rng = np.random.Generator(np.random.PCG64(1))
for (lo, hi, st, n) in ((4.0, 8.0, 3.0, 260), (4.0, 8.0, 3.0, 30), (4.0, 10.0, 3.0, 260), (2.0, 6.0, 3.0, 1500)):
    vals = (lo + st * np.floor(rng.uniform(0, np.floor((hi - lo) / st) + 1, n)))[:, None]
    loss = rng.normal(size=n)
    cand = np.arange(np.ceil((lo - st / 2) / st) * st, hi + st / 2, st)[:, None].astype(float)
    cand = np.concatenate([cand, cand + 0.25, vals[:4]])
    ctx = kb.Context(0); ctx.space_set([kb.KB_STEP_INT], [lo], [hi], [st], [0]); ctx.tpe_set_history(vals, loss)
    got = ctx.tpe_scores(cand)[:, 0]
    ref, _ = O.tpe_score_numeric(cand[:, 0], vals[:, 0], loss, lo - st / 2, hi + st / 2, st)
    pa = [ctx.tpe_get_parzen(0, s)[3] for s in (0, 1)]
    print(f"[{lo},{hi}] step {st} n={n} p_accept {pa}")
    for x, g, r in zip(cand[:, 0], got, ref): print(f"   x={x:6.2f} got {g: .6f} ref {r: .6f} {'<<<' if abs(g-r)>1e-4 else ''}")
    ctx.close()
