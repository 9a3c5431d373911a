import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as _orc
    _orc.build()
    return _orc


@pytest.fixture(scope="session")
def host_core():
    """Device emulator core compiled for the host (tests/tools/host_core.cpp) — test aid."""
    import ctypes as C
    import subprocess
    d = os.path.join(ROOT, "tests", "tools")
    so = os.path.join(d, "libhost_core.so")
    gen = os.path.join(ROOT, "softtreemax_b200", "csrc", "gen")
    srcs = [os.path.join(d, "host_core.cpp"), os.path.join(ROOT, "softtreemax_b200", "csrc", "emu_core.cuh")]
    srcs += [os.path.join(gen, f) for f in sorted(os.listdir(gen)) if f.endswith(".cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-Wno-unknown-pragmas",
                               "-o", so, srcs[0]])
    return C.CDLL(so)


def random_rom_game(orc, seed):
    """A cartridge of random bytes (valid reset vectors only): runs every opcode, documented or not,
    against TIA/RIOT at random — the widest net for bit-exact parity."""
    import numpy as np
    rng = np.random.default_rng(100 + seed)
    size = [2048, 4096, 8192][seed % 3]
    rom = rng.integers(0, 256, size=size, dtype=np.uint8)
    rom[-4], rom[-3] = 0x00, 0xF0
    if size == 8192:
        rom[4096 - 4], rom[4096 - 3] = 0x00, 0xF0
    d = dict(bank_scheme=1 if size == 8192 else 0, controller=seed % 2, actions=list(range(18))[: 3 + seed % 5],
             score_addr=[10, 76, 77], score2_addr=[255, 255, 33], lives_addr=57, lives_shift=seed % 3,
             lives_mask=7, term_addr=58, term_mask=0x81, term_value=0x80)
    return d, rom.tobytes()
