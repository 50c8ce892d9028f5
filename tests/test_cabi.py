"""CPU-side checks of the drop-in boundary: the library builds, loads, and exports every symbol
include/katib_b200.h declares; without a GPU it refuses to create a context (no CPU fallback)."""
import ctypes
import os
import subprocess

import pytest

import katib_b200 as kb


@pytest.fixture(scope="module")
def so():
    kb.build()
    return kb.lib_path()


def test_exports_every_declared_symbol(so):
    syms = kb.header_symbols()
    assert len(syms) >= 25
    out = subprocess.check_output(["nm", "-D", "--defined-only", so], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    missing = [s for s in syms if s not in exported]
    assert not missing, missing
    lib = ctypes.CDLL(so)
    for s in syms:
        assert getattr(lib, s) is not None


def test_no_torch_or_python_in_the_product(so):
    out = subprocess.check_output(["ldd", so], text=True)
    assert "torch" not in out and "python" not in out


def test_sm100a_sass_with_tma_and_mufu(so):
    sass = subprocess.check_output(["cuobjdump", "-sass", so], text=True)
    assert "sm_100a" in sass
    assert "UBLKCP" in sass      # cp.async.bulk (TMA) staging of the Parzen tables
    assert "MUFU.EX2" in sass    # the SFU-bound inner loop


def test_product_never_touches_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for sub in ("katib_b200", "include"):
        for dp, _, files in os.walk(os.path.join(root, sub)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", "Makefile")):
                    text = open(os.path.join(dp, f), errors="ignore").read()
                    assert "katib_oracle" not in text and "oracle/" not in text.replace("the CPU oracle (oracle/tpe.c)", ""), f


def test_create_fails_loudly_without_gpu(so):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(kb.KbError) as e:
        kb.Context(0)
    assert "no CPU path" in str(e.value)


def test_missing_nccl_is_an_error_message_not_a_crash(so):
    """kb_comm_unique_id needs no GPU, only libnccl: with a library name that cannot be loaded it must return
    KB_ERR_COMM with the loader's message (it used to read dlerror() twice and build a std::string from NULL)."""
    import sys
    code = ("import ctypes, sys; L = ctypes.CDLL(%r); buf = ctypes.create_string_buffer(128); "
            "rc = L.kb_comm_unique_id(buf); L.kb_last_error.restype = ctypes.c_char_p; "
            "print(rc, L.kb_last_error(None).decode())") % so
    env = dict(os.environ, KB_NCCL_LIBRARY="libnccl_that_does_not_exist.so.9")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=120)
    assert out.returncode == 0, out.stderr[-500:]
    rc, msg = out.stdout.split(None, 1)
    assert int(rc) == -4 and "libnccl_that_does_not_exist" in msg
