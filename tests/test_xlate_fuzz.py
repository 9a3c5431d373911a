"""The block translator (tools/xlate_rom.cpp) against the interpreter it is derived from, on the CPU: random-byte cartridges
(every opcode, indexed and indirect operands that land on the F8 bank-switch hotspots, wild TIA/RIOT traffic), many random
RAM images each, state and drawn frame compared after every frame.  The same translation units run on the GPU as plug-ins
(tests/test_gpu_parity.py::test_translation_plugin_bitexact)."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import ROOT, random_rom_game

pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")


def _build(tmp, rom, f8, name, cap=None):
    exe = os.path.join(tmp, "xlate_rom")
    if not os.path.exists(exe):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tools", "xlate_rom.cpp")])
    img = os.path.join(tmp, "image.bin")
    open(img, "wb").write(rom)
    subprocess.check_call([exe, img, name, "1" if f8 else "0", os.path.join(tmp, f"xl_{name}.cuh")] + ([str(cap)] if cap else []),
                          stderr=subprocess.DEVNULL)
    so = os.path.join(tmp, f"libxl_{name}.so")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-ffp-contract=off", "-I", tmp,
                           f'-DXL_HEADER="xl_{name}.cuh"', f"-DXL_CORE=Xl_{name}", "-o", so,
                           os.path.join(ROOT, "tests", "tools", "xl_fuzz.cpp")])
    return C.CDLL(so)


@pytest.mark.parametrize("seed,cap", [(11, None), (12, None), (13, None), (14, None), (15, None), (17, None), (11, 60), (14, 25)])
def test_translation_equals_interpreter_on_random_cartridges(orc, tmp_path, seed, cap):
    """cap: the translator stops after that many instructions (what translate.py does to bound nvcc's time); the rest of
    the cartridge is interpreted, through the same dispatcher."""
    d, rom = random_rom_game(orc, seed)
    lib = _build(str(tmp_path), rom, d["bank_scheme"] == 1, f"fuzz{seed}c{cap or 0}", cap)
    g = orc.Game(desc=orc.make_desc(d, len(rom)), rom=rom)
    n = 96
    rng = np.random.default_rng(1000 + seed)
    states = np.repeat(g.reset()[None], n, axis=0).copy()
    states[:, 0:128] = rng.integers(0, 256, (n, 128))           # random RAM: envs take different paths from the first instruction
    actions = np.asarray(d["actions"], np.uint8)[rng.integers(0, len(d["actions"]), n)]
    romb = np.frombuffer(rom, np.uint8).copy()
    lib.xf_run.restype = C.c_int
    rc = lib.xf_run(C.byref(g.desc), romb.ctypes.data_as(C.c_void_p), states.ctypes.data_as(C.c_void_p), n, 4,
                    actions.ctypes.data_as(C.c_void_p))
    assert rc != -2, "harness built for another image"
    assert rc == -1, f"translation and interpreter disagree: env {rc // 1000}, frame {rc % 1000}"
