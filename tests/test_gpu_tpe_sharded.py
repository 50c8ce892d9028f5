"""Multi-rank TPE on hardware: the sharded winners (ncclAllGather + tpe_merge_kernel) against the single-GPU winners and
the oracle -- tests/mp_tpe_sharded.py under torch.distributed.run, 2 ranks (skipped on a 1-GPU box; the round's log of
the 2-GPU run is profiles/r02_mp_tpe_sharded.log).  Plus the stream-ordering contract of the *_device entry points."""
import os
import subprocess
import sys

import numpy as np
import pytest

import katib_b200 as kb
from test_gpu_tpe import oracle_scores

pytestmark = pytest.mark.gpu


def test_sharded_winners_equal_single_gpu_and_oracle():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29537", os.path.join(here, "mp_tpe_sharded.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "TPE SHARDED OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


def test_device_entry_point_on_a_caller_stream_is_ordered_after_the_history_build():
    """kb_tpe_set_history returns while its kernels are still queued on the context's own (non-blocking) stream.  A
    kb_tpe_score_select_device on ANOTHER stream right after it must still see the finished tables: alternate two
    histories and expect each call's winners to be those of the history just set."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.Generator(np.random.PCG64(31))
    D, n, nc = 24, 6000, 4096
    kinds, lows, highs = [kb.KB_UNIFORM] * D, [-2.0] * D, [3.0] * D
    hist = []
    for h in range(2):
        vals = rng.uniform(-2, 3, (n, D))
        loss = np.sum((vals - (0.5 if h else -0.5)) ** 2, axis=1) + rng.normal(0, 0.3, n)
        hist.append((vals, loss))
    cand = rng.uniform(-2, 3, (nc, D))
    want = [oracle_scores(kinds, lows, highs, [0.0] * D, [0] * D, v, l, cand)[1] for v, l in hist]
    assert want[0] != want[1]
    ctx = kb.Context(0)
    ctx.space_set(kinds, lows, highs)
    t = torch.from_numpy(cand).to(dev)
    side = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    for it in range(8):
        h = it % 2
        ctx.tpe_set_history(*hist[h])
        idx = ctx.tpe_score_select_device(nc, t.data_ptr(), 0, stream=side.cuda_stream)[0]
        assert list(idx) == want[h], (it, h)
    ctx.close()
