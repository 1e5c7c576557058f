"""Commit-path MSMs over the RESIDENT generators (halo_msm_batch_dev with no base pointer): the fixed-base table path
against the ordinary pipeline (tuning key 14 = 1), by batch size.  usage: tools/commit_time.py [log_n]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200 as hb
from halo_b200._lib import lib, check

hb.init()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n = 1 << lg
u64p = ctypes.POINTER(ctypes.c_uint64)
gs = torch.empty((n, 8), dtype=torch.int64, device="cuda")
check(lib.halo_gen_points_dev(0, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n), ctypes.c_void_p(gs.data_ptr())))
sh = np.zeros((2, 8), dtype=np.uint64)
check(lib.halo_gen_points(0, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
hb.pp.set_pp(0, gs.cpu().numpy().view(np.uint64), sh[0], sh[1])
s = torch.cuda.Stream(); torch.cuda.set_stream(s)
check(lib.halo_set_stream(ctypes.c_void_p(s.cuda_stream)))
for b in (1, 3, 8, 16):
    cols = torch.randint(-(1 << 63), (1 << 63) - 1, (b * n, 4), dtype=torch.int64, device="cuda"); cols[:, 3] &= (1 << 62) - 1
    res = {}
    for off in (0, 1):
        lib.halo_set_tuning(14, off)
        out = np.zeros((b, 12), dtype=np.uint64)
        f = lambda: check(lib.halo_msm_batch_dev(0, None, ctypes.c_size_t(0), ctypes.c_void_p(cols.data_ptr()), ctypes.c_size_t(n), ctypes.c_size_t(b), out.ctypes.data_as(u64p)))
        f(); f()
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s); f(); e1.record(s); e1.synchronize(); best = min(best, e0.elapsed_time(e1))
        lib.halo_profile_enable(1); f()
        ms = (ctypes.c_float * 16)(); names = (ctypes.c_char_p * 16)()
        ns = lib.halo_profile_read(ms, names, 16)
        lib.halo_profile_enable(0)
        res[off] = (best, "  ".join(f"{names[i].decode()} {ms[i]:.2f}" for i in range(ns)), out.copy())
    lib.halo_set_tuning(14, 0)
    print(f"2^{lg} x {b}: fixed-base table {res[0][0]:.2f} ms [{res[0][1]}]\n          ordinary {res[1][0]:.2f} ms [{res[1][1]}]", flush=True)
