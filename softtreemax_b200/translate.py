"""Block translation of a cartridge image the library was not built with (the bundled ones are translated when the library
is built: build.py).  `build_plugin(rom, bank_scheme)` runs tools/xlate_rom.cpp on the image and compiles the resulting
translation unit — the same one build.py generates for a bundled cartridge — into

    $STM_XL_PLUGIN_DIR/libstm_xl_<fnv1a of the image>.so      (default: softtreemax_b200/xl_plugins/)

which `stm_ctx_create` loads when it meets that image (csrc/stm_kernels.cu find_plugin).  Needs g++ and nvcc (about a minute
per cartridge); without a plug-in an unknown image runs on the interpreter kernel.  The translation only changes HOW the
6507 instruction stream is executed (straight-line blocks of the interpreter's own exec_body): results are the interpreter's.
"""
import os
import shutil
import subprocess
import tempfile

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def rom_hash(rom):
    """FNV-1a over the image, as emu_core.cuh rom_hash and tools/xlate_rom.cpp."""
    h = 2166136261
    for b in bytes(rom):
        h = ((h ^ b) * 16777619) & 0xFFFFFFFF
    return h


def plugin_dir():
    return os.environ.get("STM_XL_PLUGIN_DIR") or os.path.join(HERE, "xl_plugins")


def plugin_path(rom):
    return os.path.join(plugin_dir(), "libstm_xl_%08x.so" % rom_hash(rom))


def build_plugin(rom, bank_scheme="none", force=False, max_instructions=400):
    """-> path of the plug-in for this image (built if missing or older than the kernel sources).
    max_instructions caps what is translated (nvcc's time grows faster than linearly with it: 220 instructions take 1.5 min,
    650 took 9); code the walk does not reach before the cap is interpreted."""
    rom = bytes(rom)
    out = plugin_path(rom)
    deps = [os.path.join(_build.CSRC, f) for f in ("emu_core.cuh", "stm_step.cuh")] + [os.path.join(ROOT, "tools", "xlate_rom.cpp")]
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    cxx = shutil.which("g++")
    if cxx is None or not os.path.exists(_build.NVCC):
        raise RuntimeError("translate.build_plugin needs g++ and nvcc")
    os.makedirs(plugin_dir(), exist_ok=True)
    name = "rom_%08x" % rom_hash(rom)
    f8 = "1" if bank_scheme in ("f8", "F8", 1) else "0"
    with tempfile.TemporaryDirectory() as tmp:
        exe = os.path.join(tmp, "xlate_rom")
        subprocess.check_call([cxx, "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tools", "xlate_rom.cpp")])
        img = os.path.join(tmp, "image.bin")
        with open(img, "wb") as f:
            f.write(rom)
        cuh = os.path.join(tmp, f"xl_{name}.cuh")
        subprocess.check_call([exe, img, name, f8, cuh, str(int(max_instructions))], stderr=subprocess.DEVNULL)
        cu = os.path.join(tmp, f"xl_{name}.cu")
        with open(cu, "w") as f:
            f.write(_build.xl_unit_source(name, os.path.join(_build.CSRC, "stm_step.cuh"), cuh, plugin=True))
        flags = [f for f in _build.FLAGS]
        part = "%s.%d.tmp" % (out, os.getpid())                 # several ranks may build the same plug-in at once
        subprocess.check_call([_build.NVCC] + flags + ["-o", part, cu])
        os.replace(part, out)
    return out
