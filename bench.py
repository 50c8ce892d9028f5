#!/usr/bin/env python
"""bench.py -- TPE candidates scored/s on BASELINE.json's headline workload.

Workload (BASELINE configs[4]): TPE, 50 parameters (20 uniform, 5 int [0,63], 5 discrete-uniform
q=0.25 on [0,16], 20 categorical with 2..16 choices), 10 000-trial history, 64 Mi candidates per GetSuggestions,
sharded over N GPUs (strong scaling: the candidate set is fixed, each rank scores 1/N of it, one NCCL collective
picks the winners).  A "step" = one full scoring+selection pass over the candidate set.

  value : candidates/s with the rank's candidate slice already resident in HBM (kb_tpe_score_select_device)
  e2e   : the same through the host-buffer C-ABI call (kb_tpe_score_select): every step copies the slice from pinned
          host memory (double-buffered H2D inside the call) and reads the winners back
  roofline : per kernel class, CUDA-event durations measured live inside the library (kb_profile); the SFU peak is
          measured in this run by kb_peak_measure (MEASURED_PEAKS.json has no MUFU figure)
  cpu_baseline : the CPU oracle (port of goptuna's algorithm, oracle/tpe.c) on a bounded candidate sample

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--cand LOG2]
  N > 1: launched by torch.distributed.run, one rank per GPU.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_HISTORY = 10_000
JSON_OUT = sys.stdout  # main() replaces it by a duplicate of the original stdout before anything native can print
# admission margins of the near-tie table (kb_tpe_score.cu: KB_MARGIN_CONT / KB_MARGIN_QUANT)
MARGIN_CONT, MARGIN_QUANT = 5e-6, 1.5e-5
CAT_CHOICES = [2, 3, 4, 6, 8, 12, 16]


def workload_space():
    import katib_b200 as kb
    kinds, lows, highs, steps, nch = [], [], [], [], []
    for i in range(20):
        s = 0.5 + 0.25 * i
        kinds.append(kb.KB_UNIFORM); lows.append(-5.0 * s); highs.append(5.0 * s); steps.append(0.0); nch.append(0)
    for i in range(5):
        kinds.append(kb.KB_INT); lows.append(0.0); highs.append(63.0); steps.append(0.0); nch.append(0)
    for i in range(5):
        kinds.append(kb.KB_DISCRETE); lows.append(0.0); highs.append(16.0); steps.append(0.25); nch.append(0)
    for i in range(20):
        kinds.append(kb.KB_CATEGORICAL); lows.append(0.0); highs.append(0.0); steps.append(0.0)
        nch.append(CAT_CHOICES[i % len(CAT_CHOICES)])
    return kinds, np.array(lows), np.array(highs), np.array(steps), nch


def workload_history(kinds, lows, highs, steps, nch, n=N_HISTORY, seed=5):
    import katib_b200 as kb
    rng = np.random.Generator(np.random.PCG64(seed))
    cols, loss = [], np.zeros(n)
    for d, k in enumerate(kinds):
        if k == kb.KB_UNIFORM:
            c = rng.uniform(lows[d], highs[d], n)
            loss += (c / (highs[d] / 5.0)) ** 2
        elif k == kb.KB_INT:
            c = np.floor(rng.uniform(lows[d], highs[d] + 1, n))
            loss += ((c - 20.0) / 16.0) ** 2
        elif k == kb.KB_DISCRETE:
            c = lows[d] + steps[d] * np.floor(rng.uniform(0, (highs[d] - lows[d]) / steps[d] + 1, n))
            loss += ((c - 5.0) / 4.0) ** 2
        else:
            c = np.floor(rng.uniform(0, nch[d], n))
            loss += 0.3 * (c != (d % nch[d]))
        cols.append(c)
    loss += rng.normal(0, 1.0, n)
    return np.stack(cols, axis=1), loss


CAND_BLOCK = 1 << 20


def device_candidates(torch, kinds, lows, highs, steps, nch, first, n, seed, device):
    """Synthetic candidates in the sampler's working domain, generated on the device by torch (plumbing):
    uniform over each parameter's range; quantised parameters on their grid; categorical = choice index.
    Rows [first, first + n) of ONE global candidate set: block b (2^20 rows) is drawn from its own generator seeded by
    (seed, b), so a row's content depends only on its global index -- every sharding of the set scores the same rows
    and must report the same winners."""
    import katib_b200 as kb
    g = torch.Generator(device=device)
    out = torch.empty((n, len(kinds)), dtype=torch.float64, device=device)
    rows = CAND_BLOCK
    for b in range(first // rows, (first + n + rows - 1) // rows):
        g.manual_seed(seed * 1_000_003 + b)
        m = rows
        u = torch.rand((m, len(kinds)), generator=g, device=device, dtype=torch.float64)
        for d, k in enumerate(kinds):
            if k == kb.KB_UNIFORM:
                u[:, d] = lows[d] + u[:, d] * (highs[d] - lows[d])
            elif k == kb.KB_INT:
                u[:, d] = torch.floor(lows[d] + u[:, d] * (highs[d] - lows[d] + 1))
            elif k == kb.KB_DISCRETE:  # working domain is shifted to [0, high-low]
                u[:, d] = steps[d] * torch.floor(u[:, d] * ((highs[d] - lows[d]) / steps[d] + 1))
            else:
                u[:, d] = torch.floor(u[:, d] * nch[d])
        g0, g1 = max(first, b * rows), min(first + n, (b + 1) * rows)  # the part of block b inside [first, first + n)
        out[g0 - first:g1 - first] = u[g0 - b * rows:g1 - b * rows]
    return out


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.p = index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                       "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        self.t.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 9 for i in range(4) if r[5 + i].lower().startswith("active")})
        pw = [float(r[3]) for r in self.rows if len(r) >= 9 and r[3].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "power_w_max": max(pw) if pw else None, "samples": len(sm)}


def oracle_score_block(O, kinds, lows, highs, steps, nch, vals, loss, cand):
    """The CPU port scoring every parameter of a candidate block (returns the per-parameter winners)."""
    import katib_b200 as kb
    best = []
    for d, k in enumerate(kinds):
        if k == kb.KB_CATEGORICAL:
            _, b = O.tpe_score_categorical(cand[:, d], vals[:, d], loss, nch[d])
        elif k == kb.KB_UNIFORM:
            _, b = O.tpe_score_numeric(cand[:, d], vals[:, d], loss, lows[d], highs[d])
        elif k == kb.KB_INT:
            _, b = O.tpe_score_numeric(cand[:, d], vals[:, d], loss, lows[d] - 0.5, highs[d] + 0.5, 1.0)
        else:
            q = steps[d]
            _, b = O.tpe_score_numeric(cand[:, d], vals[:, d] - lows[d], loss, -0.5 * q, (highs[d] - lows[d]) + 0.5 * q, q)
        best.append(b)
    return best


def cpu_baseline(kinds, lows, highs, steps, nch, vals, loss, n_sample, threads, seed=77):
    """candidates/s of the CPU port (oracle/tpe.c) on n_sample candidates, split over `threads` host threads."""
    import _oracle as O
    O.lib()
    import katib_b200 as kb
    rng = np.random.Generator(np.random.PCG64(seed))
    cols = []
    for d, k in enumerate(kinds):
        u = rng.random(n_sample)
        if k == kb.KB_UNIFORM:
            cols.append(lows[d] + u * (highs[d] - lows[d]))
        elif k == kb.KB_INT:
            cols.append(np.floor(lows[d] + u * (highs[d] - lows[d] + 1)))
        elif k == kb.KB_DISCRETE:
            cols.append(steps[d] * np.floor(u * ((highs[d] - lows[d]) / steps[d] + 1)))
        else:
            cols.append(np.floor(u * nch[d]))
    cand = np.stack(cols, axis=1)
    parts = np.array_split(np.arange(n_sample), threads)
    t0 = time.perf_counter()
    if threads == 1:
        oracle_score_block(O, kinds, lows, highs, steps, nch, vals, loss, cand)
    else:
        ths = [threading.Thread(target=oracle_score_block, args=(O, kinds, lows, highs, steps, nch, vals, loss, cand[p]))
               for p in parts if len(p)]
        [t.start() for t in ths]
        [t.join() for t in ths]
    dt = time.perf_counter() - t0
    return n_sample / dt, dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--cand", type=int, default=26, help="log2 of the candidate count (26 = 64 Mi, the headline)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="candidates in the CPU-baseline sample (0 = auto)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--workload", default="tpe50", help="tpe50 (headline, configs[4]) | sobol | tpe20 | cmaes | tpe3")
    args = ap.parse_args()
    if args.workload != "tpe50":
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_workloads
        if int(os.environ.get("RANK", "0")) == 0:
            bench_workloads.WORKLOADS[args.workload](args)
        return
    # NCCL reports its communicators (rank count, transport) at level INFO on the process's stdout (fd 1).  stdout must
    # carry exactly one JSON line, so fd 1 is pointed at stderr for everything native and the JSON line goes to a
    # duplicate of the original stdout (JSON_OUT)
    global JSON_OUT
    sys.stdout.flush()
    JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    os.environ["NCCL_DEBUG"] = os.environ.get("KB_NCCL_DEBUG", "INFO")  # the box presets a quieter level
    os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n_total = 1 << args.cand
    kinds, lows, highs, steps, nch = workload_space()
    vals, loss = workload_history(kinds, lows, highs, steps, nch)
    D = len(kinds)
    n_below = min(int(np.ceil(0.1 * N_HISTORY)), 25)
    K_l, K_g = n_below + 1, N_HISTORY - n_below + 1
    config = {"workload": "TPE 50-dim mixed (20 uniform, 5 int[0,63], 5 discrete q=0.25 on [0,16], 20 categorical), "
                          f"{N_HISTORY}-trial history, 2^{args.cand} candidates per GetSuggestions, sharded across N GPUs",
              "baseline_config": "configs[4]", "n_candidates": n_total, "n_params": D, "n_trials": N_HISTORY,
              "parallelism": f"candidate-sharded x{world}", "l2": "inputs larger than L2 (per-step candidate slice >= 400 MB)"}

    # ------------------------------------------------------------------ reference arm: CPU port on host cores
    if args.impl == "reference":
        if rank != 0:
            return
        cores = os.cpu_count() or 1
        n_sample = args.cpu_sample or 256 * cores
        for _ in range(max(1, args.warmup)):
            cpu_baseline(kinds, lows, highs, steps, nch, vals, loss, max(cores, n_sample // 8), cores)
        rates, t_tot = [], 0.0
        for _ in range(args.steps):
            r, dt = cpu_baseline(kinds, lows, highs, steps, nch, vals, loss, n_sample, cores)
            rates.append(r); t_tot += dt
        v = n_sample * args.steps / t_tot
        print(json.dumps({
            "impl": "reference", "metric": "tpe_candidates_scored_per_s", "value": v, "unit": "candidates/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config,
            "cpu_baseline": {"value": v, "unit": "candidates/s", "cores": cores, "kind": "port",
                             "sample": f"{n_sample} candidates x {D} params x {N_HISTORY}-trial history per step "
                                       "(goptuna is Go and un-vendored: the oracle port stands in; it scores one "
                                       "candidate block per host thread)"},
            "e2e": {"value": v, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}),
            file=JSON_OUT, flush=True)
        return

    # ------------------------------------------------------------------ our arm
    import torch
    import katib_b200 as kb
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    ctx = kb.Context(local_rank)
    if world > 1:
        uid = [kb.Context.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(world, rank, uid[0])
    ctx.space_set(kinds, lows, highs, steps, nch)
    ctx.tpe_set_history(vals, loss)
    from katib_b200.sharding import shard_range
    offset, end = shard_range(n_total, world, rank)
    n_local = end - offset
    cand = device_candidates(torch, kinds, lows, highs, steps, nch, offset, n_local, 55, dev)
    torch.cuda.synchronize()
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            res = fn()
        e1.record(stream)
        e1.synchronize()
        barrier()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms, wall * 1e3], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1]) / 1e3
        return ms, wall, res

    step_dev = lambda: ctx.tpe_score_select_device(n_local, cand.data_ptr(), offset)
    peaks = {}
    if rank == 0:
        peaks = {"mufu_ex2_gops": ctx.peak(0), "fp32_gfma": ctx.peak(1), "fp64_gfma": ctx.peak(2), "hbm_write_gbs": ctx.peak(3),
                 "fp32_gfma_ffma2_two_varying_operands": ctx.peak(7)}
    for _ in range(args.warmup):
        step_dev()
    launches0 = ctx.launch_count()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ctx.profile(True)
    ms, wall, res = timed(step_dev, args.steps)
    prof = {k: ctx.profile_read(k) for k in range(4)}
    ctx.profile(False)
    clk = clocks.stop() if rank == 0 else None
    launches = ctx.launch_count() - launches0
    value = n_total * args.steps / (ms * 1e-3)
    own, wide = ctx.tpe_local_stats()  # warm-up + timed steps of this rank
    local_blocks = {"own_origin": own, "hi_lo_form": wide,
                    "note": "thread blocks of tpe_score_cont_local_kernel (KB_OPT_LOCAL_ORIGIN, default on): candidates bucketed by "
                            "value, x - mu measured from the block's own origin (3 instead of 5 packed instructions per pair-component); "
                            "blocks whose candidates span more than 1/128 of the range keep the (hi, lo) form"}

    # ---- end to end: host buffers through the public C-ABI call
    e2e = None
    if not args.no_e2e:
        try:
            host = torch.empty((n_local, D), dtype=torch.float64, pin_memory=True)
            host.copy_(cand)
            pinned_note = "full slice pinned"
        except RuntimeError as ex:  # not enough pinnable memory on this box
            host, pinned_note = None, f"pinned allocation failed: {str(ex)[:80]}"
        if host is not None:
            torch.cuda.synchronize()
            step_host = lambda: ctx.tpe_score_select_ptr(n_local, host.data_ptr(), offset)
            for _ in range(args.warmup):
                step_host()
            ms_h, wall_h, res_h = timed(step_host, args.steps)
            assert all(np.array_equal(a, b) for a, b in zip(res, res_h)), "e2e winners differ from the device-resident run"
            e2e = {"value": n_total * args.steps / wall_h, "unit": "candidates/s",
                   "h2d_bytes_per_step": int(n_local * D * 8 * world), "d2h_bytes_per_step": int(D * 24 * world),
                   "ms_per_step": 1e3 * wall_h / args.steps, "timing": "host wall clock around the blocking C-ABI call, "
                   "max over ranks", "note": pinned_note}
        else:
            e2e = {"value": None, "unit": "candidates/s", "note": pinned_note}

    # ---- secondary: the whole of tpe.Sampler.Sample at this size -- kb_tpe_suggest draws its 2^cand candidates from l(x)
    # on the device (chunk by chunk, straight into the staged columns: no candidate matrix), scores and selects; with
    # N ranks every rank draws and scores its index range of the one stream.  Against `value` this shows what the
    # sampling costs.
    suggest = None
    if os.environ.get("KB_BENCH_SUGGEST", "1") == "1":
        step_sug = lambda: ctx.tpe_suggest(n_total, 4242)
        for _ in range(min(2, args.warmup)):
            out_sug = step_sug()
        ms_s, wall_s, out_sug = timed(step_sug, args.steps)
        scratch = None
        if world == 1:  # what a context that only ever suggests holds on the device: measured on a fresh one
            torch.cuda.synchronize()
            f0 = torch.cuda.mem_get_info(dev)[0]
            c2 = kb.Context(local_rank)
            c2.space_set(kinds, lows, highs, steps, nch)
            c2.tpe_set_history(vals, loss)
            same = bool(np.array_equal(c2.tpe_suggest(n_total, 4242), out_sug))
            scratch = {"bytes": int(f0 - torch.cuda.mem_get_info(dev)[0]), "same_suggestion": same,
                       "note": "device memory a fresh context takes for history, tables, staging of one chunk and near-tie tables"}
            c2.close()
        suggest = {"value": n_total * args.steps / (ms_s * 1e-3), "unit": "candidates/s (drawn, scored, selected)",
                   "ms_per_step": ms_s / args.steps, "fresh_context_device_memory": scratch,
                   "suggestion_head": [float(x) for x in out_sug[:4]],
                   "note": "kb_tpe_suggest(n_ei_candidates = the whole set): Philox draw from the below mixture fused into the "
                           "staging kernel, same scoring kernels; the candidates are never stored"}

    # ---- secondary: the same step with KB_OPT_MERGE_DUPLICATES (what kb_service_* runs with).  Reported beside the
    # headline, never as it: the headline and the roofline count every one of the reference's n+1 components.
    merged = None
    if os.environ.get("KB_BENCH_MERGED", "1") == "1":
        ctx.tpe_set_option(1, 1)
        ctx.tpe_set_history(vals, loss)
        k_merged = [len(ctx.tpe_get_parzen(d, 1)[0]) for d in range(D)] if rank == 0 else None
        for _ in range(min(2, args.warmup)):
            step_dev()
        ms_m, wall_m, res_m = timed(step_dev, args.steps)
        assert np.array_equal(res[0], res_m[0]), "merged-table winners differ from the unmerged run"
        merged = {"value": n_total * args.steps / (ms_m * 1e-3), "unit": "candidates/s", "ms_per_step": ms_m / args.steps,
                  "components_above_per_param": k_merged,
                  "note": "identical mixture and winners; duplicate (mu, sigma) components of int/discrete parameters folded"}
        ctx.tpe_set_option(1, 0)
    # ---- secondary: KB_OPT_PRUNE_FAR ("also report the pruned variant, roofline always from the
    # algorithmic count"), alone and together with the merged tables; winners must not move
    pruned = None
    if os.environ.get("KB_BENCH_PRUNED", "1") == "1":
        pruned = {}
        for name, merge in (("prune_far", 0), ("prune_far+merge_duplicates", 1)):
            ctx.tpe_set_option(1, merge)
            ctx.tpe_set_option(2, 1)
            for _ in range(min(2, args.warmup)):
                step_dev()
            ms_p, wall_p, res_p = timed(step_dev, args.steps)
            assert np.array_equal(res[0], res_p[0]), f"{name}: winners differ from the unpruned run"
            pruned[name] = {"value": n_total * args.steps / (ms_p * 1e-3), "unit": "candidates/s", "ms_per_step": ms_p / args.steps}
        pruned["note"] = ("continuous parameters skip (candidate, component) pairs more than 7.75 floor-sigmas apart "
                          "(< e^-30 of a component's peak); identical winners")
        ctx.tpe_set_option(2, 0)
        ctx.tpe_set_option(1, 0)

    # ---- parity inside the bench (untimed): winners of a prefix of THIS candidate set against the CPU oracle, and the
    # worst FP32-pipeline score error against the FP64 decision formula evaluated on the device for every pair
    winners_vs_oracle, score_error = None, None
    if rank == 0 and os.environ.get("KB_BENCH_PARITY", "1") == "1":
        import _oracle as O
        O.lib()
        cores = os.cpu_count() or 1
        n_pref = int(os.environ.get("KB_BENCH_PARITY_ROWS", str(65536 if cores >= 16 else 16384)))
        n_pref = min(n_pref, n_local)
        pref = cand[:n_pref].cpu().numpy()
        pctx = kb.Context(local_rank)  # no communicator: rank 0 alone does this
        pctx.space_set(kinds, lows, highs, steps, nch)
        pctx.tpe_set_history(vals, loss)
        got_idx = pctx.tpe_score_select_device(n_pref, cand.data_ptr(), offset)[0]
        parts = np.array_split(np.arange(n_pref), cores)
        scores = [None] * len(parts)

        def work(k):
            sc = np.empty((len(parts[k]), D))
            for d, kd in enumerate(kinds):
                if kd == kb.KB_CATEGORICAL:
                    sc[:, d] = O.tpe_score_categorical(pref[parts[k], d], vals[:, d], loss, nch[d])[0]
                elif kd == kb.KB_UNIFORM:
                    sc[:, d] = O.tpe_score_numeric(pref[parts[k], d], vals[:, d], loss, lows[d], highs[d])[0]
                elif kd == kb.KB_INT:
                    sc[:, d] = O.tpe_score_numeric(pref[parts[k], d], vals[:, d], loss, lows[d] - 0.5, highs[d] + 0.5, 1.0)[0]
                else:
                    q = steps[d]
                    sc[:, d] = O.tpe_score_numeric(pref[parts[k], d], vals[:, d] - lows[d], loss, -0.5 * q,
                                                   (highs[d] - lows[d]) + 0.5 * q, q)[0]
            scores[k] = sc
        t0 = time.perf_counter()
        ths = [threading.Thread(target=work, args=(k,)) for k in range(len(parts)) if len(parts[k])]
        [t.start() for t in ths]
        [t.join() for t in ths]
        sc_all = np.concatenate([s for s in scores if s is not None], axis=0)
        want_idx = [int(np.argmax(sc_all[:, d])) + offset for d in range(D)]  # argmax: first maximum
        winners_vs_oracle = {"n": n_pref, "identical": bool([int(i) for i in got_idx] == want_idx),
                             "oracle_s": time.perf_counter() - t0, "oracle_threads": len(ths),
                             "note": "winners of the first n rows of this run's candidate set: kb_tpe_score_select_device vs "
                                     "oracle/tpe.c (first maximum of the FP64 scores)"}
        n_err = min(n_local, 1 << int(os.environ.get("KB_BENCH_ERR_LOG2", "21")))
        t0 = time.perf_counter()
        errs = pctx.tpe_score_error_device(n_err, cand.data_ptr())
        pctx.close()
        e_cont = float(max(errs[d] for d, kd in enumerate(kinds) if kd == kb.KB_UNIFORM))
        e_quant = float(max(errs[d] for d, kd in enumerate(kinds) if kd in (kb.KB_INT, kb.KB_DISCRETE, kb.KB_STEP_INT)))
        score_error = {"rows": n_err, "pairs": n_err * 30, "continuous": e_cont, "quantised": e_quant,
                       "margin_continuous": MARGIN_CONT, "margin_quantised": MARGIN_QUANT,
                       "margin_over_error": [MARGIN_CONT / max(e_cont, 1e-300), MARGIN_QUANT / max(e_quant, 1e-300)],
                       "seconds": time.perf_counter() - t0,
                       "note": "max |score32 - score64| / max(1, |score64|) over every (row, numeric parameter) pair of the first "
                               "`rows` candidates, FP64 side = the decision formula on the device (kb_tpe_score_error_device); "
                               "KB_BENCH_ERR_LOG2=26 covers the whole set (minutes) -- that run is kept in profiles/"}

    # ---- secondary: BASELINE configs[1], [2], [3] on this GPU (rank 0), compact
    secondary = None
    if rank == 0 and os.environ.get("KB_BENCH_SECONDARY", "1") == "1":
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_workloads
        host = None  # release the pinned slice before the Sobol block needs its own host array
        secondary = bench_workloads.secondary_block(args, local_rank)

    def shutdown():
        # orderly exit on every rank: the library's NCCL communicator goes first, on all ranks together, then torch's
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        ctx.close()
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()

    if rank != 0:
        shutdown()
        return
    # ---- roofline from the live per-kernel event times (this rank's slice; every rank does the same work)
    n_cont, n_quant = 20, 10
    launch_rows = min(n_local, 1 << int(os.environ.get("KB_CHUNK_LOG2", "22")))  # candidates per kernel launch (staged chunk)
    ce_cont = float(n_local) * n_cont * (K_l + K_g) * args.steps
    ce_quant = float(n_local) * n_quant * (K_l + K_g) * args.steps
    names = ["tpe_stage_kernel", "tpe_score_cont_local_kernel (+ tpe_bucket_count / _scan / _scatter, timed with it)", "tpe_score_quant_grid_kernel (+ tpe_score_quant_kernel for off-grid tiles: none here)",
             "tpe_rescore+pick"]
    tot = sum(p[0] for p in prof.values()) or 1.0
    kern = []
    for k in range(4):
        ms_k, n_k = prof[k]
        ent = {"name": names[k], "ms_total": ms_k, "launches": n_k, "share_of_kernel_time": ms_k / tot}
        if k == 1 and ms_k > 0:
            a = ce_cont / (ms_k * 1e-3) / 1e9
            # the kernel sends 48 of every 64 exponentials through MUFU.EX2 (56 of 64 in the ordered kernel) and computes
            # the rest on the FMA pipe: its SFU ceiling is the measured MUFU.EX2 peak times 64/48 (64/56)
            on_fma = 16 if own + wide > 0 else 8
            sfu_ceiling = peaks["mufu_ex2_gops"] * 64.0 / (64 - on_fma)
            ent.update(bound="sfu", unit="G component-evals/s (one exponential each)", achieved=a, peak=sfu_ceiling,
                       frac=a / sfu_ceiling, mufu_ex2_peak=peaks["mufu_ex2_gops"], frac_of_mufu_ex2_peak=a / peaks["mufu_ex2_gops"],
                       traffic=273.9e6 * launch_rows / float(1 << 20),
                       note=f"{on_fma} of every 64 exponentials (12 of 64 for the few wide components) are computed on the FMA pipe "
                            f"(packed FP32 polynomial), the other {64 - on_fma} by MUFU.EX2: `peak` is the measured MUFU.EX2 peak "
                            f"(`mufu_ex2_peak`) times 64/{64 - on_fma}, the rate at which the XU pipe would be saturated; "
                            "`frac_of_mufu_ex2_peak` is the same rate against the MUFU.EX2 peak alone (the figure of earlier "
                            "rounds).  Every (candidate, component) pair is evaluated.  The time includes the three bucketing "
                            "kernels that put the candidates in value order for it",
                       traffic_note=f"dram read+write bytes per launch: 273.9 MB measured by ncu --set full on a 2^20-candidate launch "
                                    f"(profiles/r02_ncu_tpe_score.md), scaled to this run's {launch_rows}-candidate launches "
                                    "(the candidate columns in bucket order, read for the block's range and once per side, "
                                    "and the index permutation)")
        if k == 2 and ms_k > 0:
            a = ce_quant / (ms_k * 1e-3) / 1e9
            ent.update(bound="fp32", unit="G bin-integral evals/s (1 FP32 FMA lane-op each as executed: look-up class)", achieved=a,
                       peak=peaks["fp32_gfma"], frac=a / peaks["fp32_gfma"], traffic=200.3e6 * launch_rows / float(1 << 20),
                       note="every (candidate, component) pair of this workload is an on-grid pair: its bin integral is a table "
                            "value times the component weight, one FP32 FMA lane-op; the bound is the FP32 FMA pipe "
                            "(kb_peak_measure FFMA micro-kernel).  Against the erf formulation's SFU roofline (2 exp per "
                            f"evaluation) the same rate is {a / (peaks['mufu_ex2_gops'] / 2):.2f}x; against an FFMA2 stream with this "
                            "kernel's operand pattern (accumulator and table value both change every instruction: "
                            f"kb_peak_measure 7) it is {a / peaks['fp32_gfma_ffma2_two_varying_operands']:.2f}",
                       traffic_note=f"dram read+write bytes per launch: 200.3 MB measured by ncu --set full on a 2^20-candidate launch "
                                    f"(profiles/r02_ncu_tpe_score.md), scaled to this run's {launch_rows}-candidate launches")
        if k == 0 and ms_k > 0:
            bytes_alg = float(n_local) * D * 8 * args.steps + float(n_local) * (n_cont * 8 + n_quant * 16) * args.steps
            a = bytes_alg / (ms_k * 1e-3) / 1e9
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
                os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
            ent.update(bound="hbm", unit="GB/s", achieved=a, peak=mp, frac=a / mp)
        kern.append(ent)
    dom = max((e for e in kern if "frac" in e), key=lambda e: e["ms_total"])
    roofline = {"bound": dom["bound"], "achieved": dom["achieved"], "peak": dom["peak"], "unit": dom["unit"],
                "frac": dom["frac"], "traffic": dom.get("traffic"), "kernel": dom["name"],
                "peak_source": "kb_peak_measure MUFU.EX2 micro-kernel on this GPU in this run (of measured), scaled by the share of "
                               "exponentials the kernel sends to MUFU.EX2 (see the kernel's note); MEASURED_PEAKS.json has HBM "
                               "and bf16 only",
                "frac_of_mufu_ex2_peak": dom.get("frac_of_mufu_ex2_peak"),
                "kernels": kern, "measured_peaks": peaks}
    # BASELINE.json's metric names two figures: candidates/s (above) and the GetSuggestions p50 latency, which is quoted on
    # configs[0] (3 parameters, 100 trials, 24 candidates) -- measured here on rank 0's GPU, a few hundred calls
    latency = None
    if os.environ.get("KB_BENCH_LATENCY", "1") == "1":
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_workloads
        lt = bench_workloads.tpe3(args, calls=300, device=local_rank, emit=False)
        latency = {k: lt[k] for k in ("metric", "value", "unit", "p95_ms", "suggest_only_p50_ms", "service_p50_ms", "service_p95_ms")}
        latency["config"] = lt["config"]
        latency["cpu_baseline_ms"] = lt["cpu_baseline"]["value"]
    cores = 1
    n_sample = args.cpu_sample or 384
    cpu_v, cpu_dt = cpu_baseline(kinds, lows, highs, steps, nch, vals, loss, n_sample, cores)
    out = {"metric": "tpe_candidates_scored_per_s", "value": value, "unit": "candidates/s", "n_gpus": world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f32 (hi/lo-split inputs, f64 accumulation; winners re-scored in f64)",
           "data": "synthetic", "config": config, "clocks": clk, "e2e": e2e, "gpu_launches": launches,
           "roofline": roofline,
           "cpu_baseline": {"value": cpu_v, "unit": "candidates/s", "cores": cores, "kind": "port",
                            "sample": f"{n_sample} candidates x {D} params x {N_HISTORY}-trial history, {cpu_dt:.1f} s, "
                                      "oracle/tpe.c single thread (goptuna's path is sequential: service.go:77-88)"},
           "suggest_whole_set": suggest, "merged_duplicates": merged, "pruned": pruned, "latency": latency,
           "winners_vs_oracle": winners_vs_oracle, "score_error_max": score_error, "secondary": secondary,
           "local_origin_blocks": local_blocks,
           "winners_index_head": [int(i) for i in res[0][:6]], "wall_s_timed_region": wall}
    print(json.dumps(out), file=JSON_OUT, flush=True)
    shutdown()


if __name__ == "__main__":
    main()
