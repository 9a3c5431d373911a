"""Per-source-line instruction cost of the K2 step kernel from an `ncu --set full --import-source on`
capture: joins ncu's per-SASS "Instructions Executed" with nvdisasm's line table of the shipped cubin.

    python tools/k2_line_profile.py gpurun_out/prof.ncu-rep [top_n]
Prints SASS instructions per emulated instruction (loop iteration), attributed to source lines.
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line_table(want_len=None):
    so = os.path.join(ROOT, "softtreemax_b200", "libstm_b200.so")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=d, capture_output=True)
        # the interpreter kernel lives in stm_kernels.cu, every translated cartridge in its own translation unit
        cubs = [f for f in os.listdir(d) if f.endswith(".cubin") and (f.startswith("stm_kernels") or f.startswith("xl_"))]
        txt = "".join(subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, c)], capture_output=True, text=True).stdout for c in cubs)
    secs, cur, cf, cl = {}, None, None, None
    for l in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
        if m:
            cur = m.group(1)
            secs[cur] = []
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            cf, cl = os.path.basename(m.group(1)), int(m.group(2))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m and cur:
            secs[cur].append((cf, cl))
    cands = [v for k, v in secs.items() if "stm_step_kernel" in k]
    if want_len is not None:                       # several instantiations (interpreter, translated cartridges): match by size
        for v in cands:
            if len(v) == want_len:
                return v
    return cands[0]


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    k = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    hdr = rows[k[0] + 1]
    ie = hdr.index("Instructions Executed")
    isamp = [i for i, h in enumerate(hdr) if h.startswith("Warp Stall Sampling (All")]
    isamp = isamp[0] if isamp else None
    first = rows[k[0] + 2:(k[1] if len(k) > 1 else len(rows))]
    table = line_table(len(first))
    if len(table) != len(first):
        sys.exit(f"SASS of the capture ({len(first)}) and of the library ({len(table)}) differ: rebuild / recapture")
    counts = collections.Counter(int(r[ie]) for r in first)
    n_iter = max(c for c, n in counts.items() if n >= 10 and c > 0)          # the loop-head count
    agg, samp = collections.Counter(), collections.Counter()
    for r, (f, ln) in zip(first, table):
        agg[(f, ln)] += int(r[ie])
        if isamp is not None:
            samp[(f, ln)] += int(r[isamp] or 0)
    tot = sum(agg.values())
    stot = max(1, sum(samp.values()))
    src = {}
    for f in ("emu_core.cuh", "stm_kernels.cu", "stm_step.cuh"):
        src[f] = open(os.path.join(ROOT, "softtreemax_b200", "csrc", f)).read().splitlines()
    gen = os.path.join(ROOT, "softtreemax_b200", "csrc", "gen")
    for f in os.listdir(gen) if os.path.isdir(gen) else []:
        src[f] = open(os.path.join(gen, f)).read().splitlines()
    print(f"SASS instructions per emulated instruction: {tot / n_iter:.1f}")
    for (f, ln), c in agg.most_common(top):
        text = src[f][ln - 1].strip()[:100] if f in src and ln - 1 < len(src[f]) else ""
        print(f"{c / n_iter:7.2f} {100.0 * samp[(f, ln)] / stot:5.1f}%  {f}:{ln:<5d} {text}")
    # coarse split by source region: the TIA renderer, the bus, the 6507 core, the kernel wrapper
    emu = src["emu_core.cuh"]
    marks = [(i + 1, l.strip()) for i, l in enumerate(emu) if re.match(r"\s*(STM_DEV|STM_FN|STM_NOINLINE|template|struct)\b", l)]
    def region(f, ln):
        if f != "emu_core.cuh":
            return f
        name = "?"
        for m, t in marks:
            if m <= ln:
                name = t[:70]
        return name
    reg, regs = collections.Counter(), collections.Counter()
    for k2, c in agg.items():
        reg[region(*k2)] += c
        regs[region(*k2)] += samp[k2]
    print("-- by function (SASS/emulated instr, % of stall samples)")
    for name, c in reg.most_common(25):
        print(f"{c / n_iter:7.2f} {100.0 * regs[name] / stot:5.1f}%  {name}")


if __name__ == "__main__":
    main()
