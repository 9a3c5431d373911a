"""GPU probe (no product path uses it): can a tcgen05 A operand start at a 128-byte-aligned, not 1024-byte-aligned,
address of a 128B-swizzled tile, and with which base-offset value?  Prints the error of every (shift, base_offset)."""
import ctypes as C
import sys
import torch
sys.path.insert(0, '.')
from softtreemax_b200._lib import lib
from softtreemax_b200.engine import Engine

eng = Engine("synth_breakout")
L = lib()
L.stm_debug_shift_probe.restype = C.c_int
L.stm_debug_shift_probe.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
torch.manual_seed(0)
rows = 160
A = torch.randn(rows, 64, device=eng.device).to(torch.bfloat16)
B = torch.randn(64, 64, device=eng.device).to(torch.bfloat16)
out = torch.zeros(128, 64, device=eng.device)
for shift in (0, 1, 2, 3, 5, 8, 9, 20):
    res = []
    for bo in range(8):
        out.zero_()
        rc = L.stm_debug_shift_probe(eng._h, A.data_ptr(), B.data_ptr(), out.data_ptr(), rows, shift, bo, None)
        torch.cuda.synchronize()
        want = A[shift:shift + 128].float() @ B.float().t()
        res.append(float((out - want).abs().max()))
    print("shift", shift, "max|err| by base_offset 0..7:", ["%.3g" % r for r in res], flush=True)
