"""The C-ABI shared library loads here (no GPU), exports every symbol include/stm_b200.h declares,
and refuses to compute without a CUDA device instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared():
    """every function declared in the product header and in the measurement-only one"""
    names = set()
    for h in ("stm_b200.h", "stm_b200_debug.h"):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b(stm_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    from softtreemax_b200 import _lib
    L = C.CDLL(_lib.LIB_PATH)
    names = declared()
    assert "stm_bfs" in names and "stm_step" in names and len(names) >= 12
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/ but not exported"
    assert set(_lib.exported_symbols()) == set(names), "ctypes signature table out of sync with the header"


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from softtreemax_b200 import _lib, games
    d, rom, _ = games.load("BreakoutNoFrameskip-v4")
    h = C.c_void_p()
    romb = (C.c_uint8 * len(rom)).from_buffer_copy(rom)
    rc = _lib.lib().stm_ctx_create(C.byref(h), C.byref(d), romb, 0)
    assert rc == -2 and b"no CPU fallback" in _lib.lib().stm_last_error(None)
    with pytest.raises(_lib.StmError):
        from softtreemax_b200.engine import Engine
        Engine("synth_breakout")


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "softtreemax_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f)).read()
                assert "from oracle" not in txt and "import oracle" not in txt and "libstm_oracle" not in txt, f


def test_state_layout_constants():
    from softtreemax_b200 import _lib
    assert C.sizeof(_lib.GameDesc) == 40
