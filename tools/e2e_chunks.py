"""End-to-end halo_msm (host buffers) at 2^20 points for different numbers of point ranges (tuning key 10), pinned and
pageable buffers.  usage: tools/e2e_chunks.py [log_n]"""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
k = 1 << lg
u64p = ctypes.POINTER(ctypes.c_uint64)
d_pts = torch.empty((k, 8), dtype=torch.int64, device="cuda")
check(lib.halo_gen_points_dev(0, ctypes.c_uint64(3), ctypes.c_size_t(0), ctypes.c_size_t(k), ctypes.c_void_p(d_pts.data_ptr())))
d_sc = torch.randint(-(1 << 63), (1 << 63) - 1, (k, 4), dtype=torch.int64, device="cuda")
d_sc[:, 3] &= (1 << 62) - 1
hp, hs = d_pts.cpu().pin_memory(), d_sc.cpu().pin_memory()
pp, ps = hp.numpy().copy(), hs.numpy().copy()
out = np.zeros(12, dtype=np.uint64)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(bp, sp):
    ts = []
    for it in range(8):
        flush.zero_(); torch.cuda.synchronize()
        t = time.perf_counter()
        check(lib.halo_msm(0, ctypes.cast(bp, u64p), ctypes.c_size_t(0), ctypes.cast(sp, u64p), ctypes.c_size_t(k), out.ctypes.data_as(u64p)))
        ts.append((time.perf_counter() - t) * 1e3)
    return min(ts[2:]), sorted(ts[2:])[len(ts[2:]) // 2]
for nch in (0, 2, 3, 4, 6, 8, 12, 16):
    lib.halo_set_tuning(10, nch)
    a = run(hp.data_ptr(), hs.data_ptr()); b = run(pp.ctypes.data, ps.ctypes.data)
    print(f"ranges {nch or 'default':>7}: pinned min {a[0]:.3f} med {a[1]:.3f} ms   pageable min {b[0]:.3f} med {b[1]:.3f} ms", flush=True)
lib.halo_set_tuning(10, 0); lib.halo_set_tuning(9, 1)
a = run(hp.data_ptr(), hs.data_ptr()); b = run(pp.ctypes.data, ps.ctypes.data)
print(f"unchunked      : pinned min {a[0]:.3f} med {a[1]:.3f} ms   pageable min {b[0]:.3f} med {b[1]:.3f} ms")
