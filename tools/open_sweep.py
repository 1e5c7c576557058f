"""IPA opening time against the fold-free switch point (tuning key 3) and the one-launch small-MSM path (key 1).
usage: tools/open_sweep.py [log_n ...]"""
import ctypes, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
u64p = ctypes.POINTER(ctypes.c_uint64)
for lg in [int(x) for x in sys.argv[1:]] or [12, 16, 20]:
    n = 1 << lg
    gs = torch.empty((n, 8), dtype=torch.int64, device="cuda")
    check(lib.halo_gen_points_dev(0, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n), ctypes.c_void_p(gs.data_ptr())))
    sh = np.zeros((2, 8), dtype=np.uint64)
    check(lib.halo_gen_points(0, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
    halo_b200.pp.reset(0)
    halo_b200.pp.set_pp(0, gs.cpu().numpy().view(np.uint64), sh[0], sh[1])
    t = torch.randint(-(1 << 63), (1 << 63) - 1, (n + 1, 4), dtype=torch.int64); t[:, 3] &= (1 << 62) - 1
    p = t.numpy().view(np.uint64)
    ref = None
    for small_off in (1, 0):
        for tail in (15, 14, 13, 12, 11):
            lib.halo_set_tuning(1, small_off); lib.halo_set_tuning(3, tail)
            best = 1e9
            for _ in range(3):
                t0 = time.perf_counter(); q = halo_b200.pcdl.Instance.open(0, p[:n], n - 1, p[n]); best = min(best, time.perf_counter() - t0)
            if ref is None: ref = q.w.copy()
            same = bool((q.w[13:21] == ref[13:21]).all())     # z, v
            print(f"n=2^{lg} small_path={'off' if small_off else 'on '} tail_lg={tail}: open {best * 1e3:.2f} ms  (v matches: {same})", flush=True)
    lib.halo_set_tuning(1, 0); lib.halo_set_tuning(3, 0)
