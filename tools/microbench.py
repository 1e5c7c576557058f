"""Measures the integer-pipe denominators on the box: raw IMAD.WIDE issue rate and Montgomery modmul/s."""
import ctypes, json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
res = {}
ms = ctypes.c_float()
out = (ctypes.c_uint64 * 4)()
for threads, bps in ((128, 4), (256, 2), (256, 4), (512, 2)):
    blocks = 148 * bps
    iters = 2000
    check(lib.halo_bench_imad(blocks, threads, iters, ctypes.byref(ms), out))
    res[f"imad_wide_per_s_{threads}x{bps}"] = blocks * threads * iters * 8 / (ms.value * 1e-3)
    for f in (0, 1):
        check(lib.halo_bench_modmul(f, blocks, threads, 500, ctypes.byref(ms), out))
        res[f"modmul_per_s_f{f}_{threads}x{bps}"] = blocks * threads * 500 * 4 / (ms.value * 1e-3)
print(json.dumps(res, indent=1))
pipes = {}
for mode, name, per in ((1, "dfma", 8), (2, "dfma+imad_wide", 16), (3, "imad32", 8), (4, "iadd_carry", 8)):
    for threads, bps in ((256, 4), (512, 2)):
        blocks = 148 * bps
        check(lib.halo_bench_pipe(mode, blocks, threads, 2000, ctypes.byref(ms), out))
        pipes[f"{name}_per_s_{threads}x{bps}"] = blocks * threads * 2000 * per / (ms.value * 1e-3)
print(json.dumps(pipes, indent=1))
