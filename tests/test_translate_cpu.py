"""Translation plug-ins, the part that needs no GPU: translate.build_plugin turns a cartridge image the library was not
built with into a shared object (tools/xlate_rom.cpp + nvcc cross-compile), named by the image's hash, exporting
`stm_xl_plugin_entry` with the library's kernel ABI.  tests/test_gpu_parity.py::test_translation_plugin_bitexact runs them."""
import ctypes as C
import os
import shutil

import pytest

from conftest import random_rom_game

pytestmark = pytest.mark.skipif(shutil.which("g++") is None or not os.path.exists("/usr/local/cuda/bin/nvcc"),
                                reason="needs g++ and nvcc")


class XlEntry(C.Structure):
    _fields_ = [("name", C.c_char_p), ("abi", C.c_uint32), ("rom_hash", C.c_uint32), ("rom_size", C.c_uint32),
                ("n_cases", C.c_uint32), ("offsets", C.c_void_p), ("launch", C.c_void_p), ("set_attrs", C.c_void_p)]


@pytest.mark.parametrize("seed", [10])          # seed 11 (8 KB, F8) is built by __graft_entry__.build() and run on the GPU
def test_plugin_builds_and_describes_its_image(seed):
    from softtreemax_b200 import translate
    d, rom = random_rom_game(None, seed)
    path = translate.build_plugin(rom, d["bank_scheme"])
    assert os.path.basename(path) == "libstm_xl_%08x.so" % translate.rom_hash(rom)
    so = C.CDLL(path)
    so.stm_xl_plugin_entry.restype = C.POINTER(XlEntry)
    e = so.stm_xl_plugin_entry().contents
    assert e.rom_hash == translate.rom_hash(rom) and e.rom_size == len(rom)
    assert e.n_cases > 0 and e.launch and e.set_attrs
    assert (e.abi >> 24) == 2, "kXlAbi version byte"
    # a second call is a cache hit
    t = os.path.getmtime(path)
    assert translate.build_plugin(rom, d["bank_scheme"]) == path and os.path.getmtime(path) == t


def test_rom_hash_is_fnv1a():
    from softtreemax_b200 import translate
    assert translate.rom_hash(b"") == 2166136261
    assert translate.rom_hash(b"a") == 0xE40C292C
