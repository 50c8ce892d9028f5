"""Secondary workloads of BASELINE.json (configs[0..3]); bench.py --workload NAME dispatches here.
Each function prints ONE JSON line in the same shape as the headline bench (metric, value, unit, roofline,
cpu_baseline, e2e).  Single GPU.  The headline (configs[4]) lives in bench.py itself."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _peaks_file():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return json.load(open(p)) if os.path.exists(p) else {"hbm_gbs": 6650.0}


def _events(torch, stream):
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


# ------------------------------------------------------------------------------------------------ configs[1]
def sobol(args, device=0, emit=True):
    """Sobol 64-dim x 2^24 points, f64, bit-exact vs the oracle; HBM-write bound (8.59 GB written)."""
    import torch
    import katib_b200 as kb
    import _oracle as O
    ctx = kb.Context(device)
    dev = torch.device("cuda", device)
    dim, n = 64, 1 << (args.cand if args.cand != 26 else 24)
    out_t = torch.empty((n, dim), dtype=torch.float64, device=dev)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    for _ in range(args.warmup):
        ctx.sobol_fill_device(dim, 0, n, out_t.data_ptr())
    torch.cuda.synchronize()
    ctx.profile(True)
    e0, e1 = _events(torch, stream)
    e0.record(stream)
    for _ in range(args.steps):
        ctx.sobol_fill_device(dim, 0, n, out_t.data_ptr())
    e1.record(stream)
    e1.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    k_ms, k_n = ctx.profile_read(4)
    ctx.profile(False)
    golden = os.path.join(ROOT, "tests", "golden")
    tail = out_t[-4096:].cpu().numpy()
    exact = bool(np.array_equal(tail, O.sobol(golden, dim, n - 4096, 4096)))
    # end to end: host buffer through kb_sobol_fill (chunked device fill, D2H double-buffered behind it); pinned
    # destination = the copies land in it directly, pageable destination = through the library's pinned staging
    del tail
    e2e = {}
    for kind in ("pinned", "pageable"):
        if kind == "pinned":
            host_t = torch.empty((n, dim), dtype=torch.float64, pin_memory=True)
            ptr, host = host_t.data_ptr(), host_t.numpy()
        else:
            host = np.empty((n, dim))
            ptr = host.ctypes.data
        assert kb.lib().kb_sobol_fill(ctx._h, dim, 0, min(n, 1 << 16), None, None, ptr) == 0  # warm-up: buffers made
        t0 = time.perf_counter()
        rc = kb.lib().kb_sobol_fill(ctx._h, dim, 0, n, None, None, ptr)
        e2e[kind] = time.perf_counter() - t0
        assert rc == 0
        exact = exact and bool(np.array_equal(host[-4096:], O.sobol(golden, dim, n - 4096, 4096)))
        exact = exact and bool(np.array_equal(host[:4096], O.sobol(golden, dim, 0, 4096)))
        del host
        host_t = None
    e2e_s = e2e["pinned"]
    # CPU port on a bounded sample
    ns = 1 << 20
    t0 = time.perf_counter()
    O.sobol(golden, dim, 0, ns)
    cpu_s = time.perf_counter() - t0
    bytes_alg = float(n) * dim * 8
    achieved = bytes_alg / (k_ms / k_n * 1e-3) / 1e9
    peak = _peaks_file()["hbm_gbs"]
    out = ({
        "metric": "sobol_points_per_s", "value": n / (ms * 1e-3), "unit": "points/s (64-dim f64)", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 state -> f64", "data": "synthetic",
        "config": {"workload": f"Sobol quasirandom 64-dim, 2^{int(np.log2(n))} points, unscaled f64, row-major", "baseline_config": "configs[1]",
                   "l2": "output (8.6 GB) larger than L2"},
        "bit_exact_vs_oracle": exact,
        "e2e": {"value": n / e2e_s, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": int(bytes_alg),
                "pageable_destination_points_per_s": n / e2e["pageable"],
                "note": "kb_sobol_fill into a pinned host array (32 MB chunks, fill and D2H overlapped, no allocation); "
                        "pageable destination: through the library's pinned staging and four host copy threads"},
        "gpu_launches": k_n,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "kernel": "sobol_fill_kernel", "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy; of measured); "
                     f"write-only ceiling measured by kb_peak_measure here: {ctx.peak(3):.0f} GB/s"},
        "cpu_baseline": {"value": ns / cpu_s, "unit": "points/s", "cores": 1, "kind": "port",
                         "sample": f"2^20 points x 64 dims, {cpu_s:.2f} s, oracle/sobol.c"}})
    ctx.close()
    del out_t
    if emit:
        print(json.dumps(out))
    return out


# ------------------------------------------------------------------------------------------------ configs[2]
def tpe20(args, device=0, emit=True):
    """TPE 20-dim continuous, 1k-trial history, 1M candidates per GetSuggestions, 1 GPU."""
    import torch
    import katib_b200 as kb
    import _oracle as O
    ctx = kb.Context(device)
    dev = torch.device("cuda", device)
    rng = np.random.Generator(np.random.PCG64(3))
    D, n, nc = 20, 1000, 1_000_000 if args.cand == 26 else 1 << args.cand
    s = np.linspace(0.5, 4.0, D)
    lows, highs = -5 * s, 5 * s
    ctx.space_set([kb.KB_UNIFORM] * D, lows, highs)
    vals = rng.uniform(lows, highs, (n, D))
    loss = np.sum((vals / s) ** 2, axis=1) + rng.normal(0, 1, n)
    ctx.tpe_set_history(vals, loss)
    cand_d = torch.empty((nc, D), dtype=torch.float64, device=dev)
    lib = kb.lib()
    assert lib.kb_tpe_sample_candidates(ctx._h, nc, 33, None, cand_d.data_ptr(), None) == 0  # drawn from l(x) on the device
    torch.cuda.synchronize()
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    step = lambda: ctx.tpe_score_select_device(nc, cand_d.data_ptr(), 0)
    mufu = ctx.peak(0)
    for _ in range(args.warmup):
        step()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # 256 MB > L2: written between timed steps
    ctx.profile(True)
    tot = 0.0
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        e0, e1 = _events(torch, stream)
        e0.record(stream)
        res = step()
        e1.record(stream)
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / args.steps
    prof = {k: ctx.profile_read(k) for k in range(4)}
    ctx.profile(False)
    host = torch.empty((nc, D), dtype=torch.float64, pin_memory=True)
    host.copy_(cand_d)
    torch.cuda.synchronize()
    ctx.tpe_score_select_ptr(nc, host.data_ptr(), 0)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res_h = ctx.tpe_score_select_ptr(nc, host.data_ptr(), 0)
    e2e_s = (time.perf_counter() - t0) / args.steps
    assert np.array_equal(res[0], res_h[0])
    # parity of the winners on a prefix the oracle can score quickly
    m = 4096
    sub = host[:m].numpy()
    idx_g = ctx.tpe_score_select(sub)[0]
    t0 = time.perf_counter()
    idx_o = [O.tpe_score_numeric(sub[:, d], vals[:, d], loss, lows[d], highs[d])[1] for d in range(D)]
    cpu_s = time.perf_counter() - t0
    ce = float(nc) * D * (26 + 976) * args.steps
    a = ce / (prof[1][0] * 1e-3) / 1e9
    out = ({
        "metric": "tpe_candidates_scored_per_s", "value": nc / (ms * 1e-3), "unit": "candidates/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 (hi/lo-split inputs; winners re-scored in f64)", "data": "synthetic",
        "config": {"workload": f"TPE 20-dim continuous, 1000-trial history, {nc} candidates per GetSuggestions, 1 GPU",
                   "baseline_config": "configs[2]", "l2": "L2 flushed (256 MB write) between timed steps"},
        "winners_identical_on_4096_prefix": bool(list(idx_g) == idx_o),
        "e2e": {"value": nc / e2e_s, "unit": "candidates/s", "h2d_bytes_per_step": nc * D * 8, "d2h_bytes_per_step": D * 24},
        "gpu_launches": sum(p[1] for p in prof.values()),
        # a 1000-trial history keeps the ordered kernel (KB_OPT_LOCAL_ORIGIN = 1 starts at 2048 trials): 56 of every 64
        # exponentials on MUFU.EX2, 8 on the FMA pipe -- the SFU ceiling is the measured MUFU.EX2 peak times 64/56
        "roofline": {"bound": "sfu", "achieved": a, "peak": mufu * 64.0 / 56.0, "unit": "G component-evals/s (one exponential each)",
                     "frac": a / (mufu * 64.0 / 56.0), "mufu_ex2_peak": mufu, "frac_of_mufu_ex2_peak": a / mufu,
                     "traffic": None, "kernel": "tpe_score_cont_kernel",
                     "kernels": [{"class": k, "ms_total": prof[k][0], "launches": prof[k][1]} for k in range(4)],
                     "peak_source": "kb_peak_measure MUFU.EX2 micro-kernel, this GPU, this run (of measured)"},
        "cpu_baseline": {"value": m / cpu_s, "unit": "candidates/s", "cores": 1, "kind": "port",
                         "sample": f"{m} candidates x 20 params x 1000-trial history, {cpu_s:.1f} s, oracle/tpe.c"}})
    ctx.close()
    if emit:
        print(json.dumps(out))
    return out


# ------------------------------------------------------------------------------------------------ configs[3]
def cmaes(args, device=0, emit=True, cpu_tell_baseline=True):
    """CMA-ES 512-dim, popsize 4096, FP64; a step = one generation (ask + tell), objective evaluated by the harness."""
    import katib_b200 as kb
    import _oracle as O
    ctx = kb.Context(device)
    rng = np.random.Generator(np.random.PCG64(4))
    n, lam = 512, 4096
    gens = args.steps if args.steps > 3 else 20
    Q, _ = np.linalg.qr(rng.normal(size=(n, n)))
    scale = np.logspace(0, 6, n)
    fobj = lambda X: np.sum((X @ Q) ** 2 * scale, axis=1)
    m0 = np.full(n, 3.0)
    ctx.cmaes_create(m0, 2.0, lam, np.full(n, -5.0), np.full(n, 5.0))
    Z = rng.normal(size=(lam, n))
    X = ctx.cmaes_ask(lam, Z); f = fobj(X)
    cpu_tell, err0 = None, None
    if cpu_tell_baseline:
        opt = O.Cma(m0, 2.0, lam)
        t0 = time.perf_counter(); opt.tell(X, f); cpu_tell = time.perf_counter() - t0   # CPU port, one tell
    ctx.cmaes_tell(X, f)
    if cpu_tell_baseline:
        err0 = float(np.max(np.abs(ctx.cmaes_state()["C"] - opt.state()[2])) / np.max(np.abs(opt.state()[2])))
    for _ in range(max(0, args.warmup - 1)):
        X = ctx.cmaes_ask(lam, rng.normal(size=(lam, n))); ctx.cmaes_tell(X, fobj(X))
    ctx.profile(True)
    t_ask = t_tell = 0.0
    import torch
    # z and x travel through pinned host memory (what the e2e contract asks for); z ~ N(0, I) is drawn by the harness
    # outside the timed calls: one draw per column, then bounds repair, so no resampling rounds
    Xp = torch.empty((lam, n), dtype=torch.float64, pin_memory=True).numpy()
    Zp = torch.empty((lam, n), dtype=torch.float64, pin_memory=True).numpy()
    for g in range(gens):
        rng.standard_normal(out=Zp)
        t0 = time.perf_counter(); X = ctx.cmaes_ask(lam, Zp, out=Xp); t1 = time.perf_counter()
        f = fobj(X)
        t2 = time.perf_counter(); ctx.cmaes_tell(X, f); t3 = time.perf_counter()
        t_ask += t1 - t0; t_tell += t3 - t2
    prof = {k: ctx.profile_read(k) for k in (7, 8, 9, 11)}
    ctx.profile(False)
    per_gen = (t_ask + t_tell) / gens
    flop_gemm = 2.0 * n * n * lam * 3 + 2.0 * n * n * n * 2   # ask, C^-1/2 Y, rank-mu; C^-1/2 and C rebuild
    eig_ms = prof[9][0] / max(1, prof[9][1])
    gemm_ms = prof[11][0] / gens  # every kb_dgemm_nt launch of a generation (ask, C^-1/2, C^-1/2 Y, rank-mu, C rebuild, warm-start G)
    flop_gemm = 2.0 * n * n * lam * 3 + 2.0 * n * n * n * 3
    dfma = ctx.peak(2)
    a = flop_gemm / 2 / (max(gemm_ms, 1e-6) * 1e-3) / 1e9
    out = ({
        "metric": "cmaes_generations_per_s", "value": 1.0 / per_gen, "unit": "generations/s (ask+tell, n=512, popsize=4096)", "n_gpus": 1,
        "steps": gens, "warmup": args.warmup, "ms_per_step": per_gen * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "CMA-ES 512-dim, popsize 4096, ask/tell with FP64 rank-mu update, rotated ellipsoid cond 1e6, z supplied by the harness, z and x through pinned host memory",
                   "baseline_config": "configs[3]", "note": "popsize is forced at the C ABI; Katib's settings cannot set it (converter.go:41-68)"},
        "first_generation_max_rel_err_C_vs_oracle": err0, "final_objective": float(fobj(ctx.cmaes_state()["mean"][None])[0]),
        "e2e": {"value": 1.0 / per_gen, "unit": "generations/s", "h2d_bytes_per_step": 2 * lam * n * 8 + lam * 8, "d2h_bytes_per_step": lam * n * 8,
                "ask_ms": 1e3 * t_ask / gens, "tell_ms": 1e3 * t_tell / gens, "note": "host wall clock around kb_cmaes_ask / kb_cmaes_tell (host buffers)"},
        "gpu_launches": int(sum(prof[k][1] for k in (7, 8, 9))), "dgemm_ms_per_generation": gemm_ms, "dgemm_launches": prof[11][1],
        "roofline": {"bound": "fp64", "achieved": a, "peak": dfma, "unit": "GFMA/s FP64 on the GEMM part (DMMA)", "frac": a / dfma, "traffic": None,
                     "kernel": "dgemm_nt_kernel", "eigh_ms_per_generation": eig_ms, "ask_kernel_ms": prof[7][0] / gens, "tell_kernel_ms": prof[8][0] / gens,
                     "peak_source": "kb_peak_measure FP64 DFMA micro-kernel, this GPU, this run (of measured)"},
        "cpu_baseline": None if cpu_tell is None else
                        {"value": 1.0 / cpu_tell, "unit": "generations/s (tell only)", "cores": 1, "kind": "port",
                         "sample": f"one Tell at n=512, popsize=4096, {cpu_tell:.1f} s, oracle/cmaes.c (cyclic Jacobi + triple loops)"}})
    ctx.close()
    if emit:
        print(json.dumps(out))
    return out


# ------------------------------------------------------------------------------------------------ configs[0]
def tpe3(args, calls=1000, device=0, emit=True):
    """GetSuggestions latency shape: 3 parameters (double, int, categorical), 100-trial history, 24 EI candidates.
    Also called by the headline bench (emit=False) for its `latency` object."""
    import katib_b200 as kb
    import _oracle as O
    ctx = kb.Context(device)
    rng = np.random.Generator(np.random.PCG64(1))
    kind = [kb.KB_UNIFORM, kb.KB_INT, kb.KB_CATEGORICAL]
    ctx.space_set(kind, [0.01, 2, 0], [0.03, 5, 0], n_choices=[0, 0, 3])
    n = 100
    vals = np.stack([rng.uniform(0.01, 0.03, n), np.floor(rng.uniform(2, 6, n)), np.floor(rng.uniform(0, 3, n))], axis=1)
    loss = (vals[:, 0] - 0.02) ** 2 * 1e4 + 0.1 * vals[:, 1] + (vals[:, 2] != 1) + rng.normal(0, 0.01, n)
    lat_full, lat_suggest = [], []
    for i in range(calls + 20):
        t0 = time.perf_counter()
        ctx.tpe_set_history(vals, loss)
        t1 = time.perf_counter()
        ctx.tpe_suggest(24, seed=10 + i)
        t2 = time.perf_counter()
        if i >= 20:
            lat_full.append(t2 - t0); lat_suggest.append(t2 - t1)
    # the same shape through the wire-level boundary: kb_service_get_suggestions on serialized api.proto messages, a
    # closed loop in which every request carries ALL trials so far (as the controller sends them) and one of them is new
    import _pb
    svc = kb.Service(device)
    space = [_pb.parameter("lr", "DOUBLE", max="0.03", min="0.01"), _pb.parameter("num-layers", "INT", max="5", min="2"),
             _pb.parameter("optimizer", "CATEGORICAL", list=["sgd", "adam", "ftrl"])]
    exp = _pb.experiment("hp", "tpe", [("random_state", "10")], "MINIMIZE", "loss", space)
    trials, lat_svc, pending = [], [], None
    for i in range(320):
        if pending is not None:
            d = dict(pending)
            fval = (float(d["lr"]) - 0.02) ** 2 * 1e4 + 0.1 * int(d["num-layers"]) + (d["optimizer"] != "adam")
            trials.append(_pb.trial(f"t{i}", pending, "SUCCEEDED", [("loss", repr(fval))]))
        req = _pb.get_suggestions_request(exp, trials, 1, i + 1)
        t0 = time.perf_counter()
        code, reply, msg = svc.get_suggestions(req)
        dt = time.perf_counter() - t0
        assert code == 0, msg
        pending = _pb.parse_reply(reply)[0]
        if len(trials) >= 100:
            lat_svc.append(dt)
    svc.close()
    cand = ctx.tpe_sample_candidates(24, seed=10)
    t0 = time.perf_counter()
    reps = 200
    for _ in range(reps):
        O.tpe_score_numeric(cand[:, 0], vals[:, 0], loss, 0.01, 0.03)
        O.tpe_score_numeric(cand[:, 1], vals[:, 1], loss, 1.5, 5.5, 1.0)
        O.tpe_score_categorical(cand[:, 2], vals[:, 2], loss, 3)
    cpu = (time.perf_counter() - t0) / reps
    p = lambda a, q: float(np.percentile(np.asarray(a) * 1e3, q))
    out = {
        "metric": "getsuggestions_latency_ms_p50", "value": p(lat_full, 50), "unit": "ms (history upload + Parzen build + sample + score + select)",
        "n_gpus": 1, "steps": calls, "warmup": 20, "ms_per_step": p(lat_full, 50), "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32/f64", "data": "synthetic",
        "config": {"workload": "TPE on the hp-tuning (lr, num-layers, optimizer) space, 100-trial history, 24 candidates", "baseline_config": "configs[0]"},
        "p95_ms": p(lat_full, 95), "suggest_only_p50_ms": p(lat_suggest, 50), "suggest_only_p95_ms": p(lat_suggest, 95),
        "service_p50_ms": p(lat_svc, 50), "service_p95_ms": p(lat_svc, 95),
        "service_note": "kb_service_get_suggestions on serialized GetSuggestionsRequest bytes, 100..320 trials resent in full every "
                        "call, one new Succeeded trial per call: protobuf parse + trial conversion + matching + history append + "
                        "Parzen build + suggest + reply encoding",
        "e2e": {"value": p(lat_full, 50), "unit": "ms", "h2d_bytes_per_step": n * 4 * 8, "d2h_bytes_per_step": 72},
        "gpu_launches": ctx.launch_count(), "roofline": None,
        "cpu_baseline": {"value": cpu * 1e3, "unit": "ms per suggestion (score+select only)", "cores": 1, "kind": "port",
                         "sample": f"{reps} repetitions, oracle/tpe.c"}}
    ctx.close()
    if emit:
        print(json.dumps(out))
    return out


def secondary_block(args, device=0):
    """BASELINE configs[1], [2], [3] for the headline line's `secondary` object: the same measurements as
    `bench.py --workload sobol|tpe20|cmaes`, a few steps each, reduced to the figures that matter."""
    import types
    a = types.SimpleNamespace(steps=3, warmup=3, cand=26)
    out = {}
    try:
        r = sobol(a, device, emit=False)
        out["sobol_2^24x64"] = {"baseline_config": "configs[1]", "points_per_s": r["value"], "ms_per_fill": r["ms_per_step"],
                                "hbm_write_gbs": r["roofline"]["achieved"], "frac_of_measured_hbm": r["roofline"]["frac"],
                                "bit_exact_vs_oracle": r["bit_exact_vs_oracle"], "e2e_points_per_s": r["e2e"]["value"],
                                "cpu_port_points_per_s_1core": r["cpu_baseline"]["value"]}
    except Exception as ex:  # a secondary figure must not take the headline line down
        out["sobol_2^24x64"] = {"error": str(ex)[:200]}
    try:
        r = tpe20(a, device, emit=False)
        out["tpe20_1k_1M"] = {"baseline_config": "configs[2]", "candidates_per_s": r["value"], "ms_per_step": r["ms_per_step"],
                              "frac_of_measured_mufu": r["roofline"]["frac_of_mufu_ex2_peak"],
                              "frac_of_sfu_ceiling": r["roofline"]["frac"], "e2e_candidates_per_s": r["e2e"]["value"],
                              "winners_identical_on_4096_prefix": r["winners_identical_on_4096_prefix"],
                              "cpu_port_candidates_per_s_1core": r["cpu_baseline"]["value"]}
    except Exception as ex:
        out["tpe20_1k_1M"] = {"error": str(ex)[:200]}
    try:
        a.steps = 20
        r = cmaes(a, device, emit=False, cpu_tell_baseline=False)
        out["cmaes_512x4096"] = {"baseline_config": "configs[3]", "generations_per_s": r["value"], "ms_per_generation": r["ms_per_step"],
                                 "eigh_ms": r["roofline"]["eigh_ms_per_generation"], "dgemm_ms": r["dgemm_ms_per_generation"],
                                 "dmma_frac_of_measured_dfma": r["roofline"]["frac"], "generations_timed": r["steps"]}
    except Exception as ex:
        out["cmaes_512x4096"] = {"error": str(ex)[:200]}
    return out


WORKLOADS = {"sobol": sobol, "tpe20": tpe20, "cmaes": cmaes, "tpe3": tpe3}
