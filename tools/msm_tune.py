"""Sweep of MSM tuning knobs at one size (device-resident inputs); prints per-stage ms."""
import ctypes, sys, os, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
k = 1 << lg
d_pts = torch.empty((k, 8), dtype=torch.int64, device="cuda")
check(lib.halo_gen_points_dev(0, ctypes.c_uint64(7), ctypes.c_size_t(0), ctypes.c_size_t(k), ctypes.c_void_p(d_pts.data_ptr())))
d_sc = torch.randint(-(1 << 63), (1 << 63) - 1, (k, 4), dtype=torch.int64, device="cuda")
d_sc[:, 3] &= (1 << 62) - 1
s = torch.cuda.Stream()          # a real stream: handle 0 would mean 'library stream' to halo_set_stream
torch.cuda.set_stream(s)
check(lib.halo_set_stream(ctypes.c_void_p(s.cuda_stream)))
check(lib.halo_profile_enable(1))
out = np.empty(12, dtype=np.uint64)
ms_buf = (ctypes.c_float * 24)(); nm_buf = (ctypes.c_char_p * 24)()
ref = None
cs = [int(x) for x in os.environ.get("CS", "15,16").split(",")]
for c, t0, t1 in itertools.product(cs, (0,), (0,)):
    lib.halo_msm_set_window(c); lib.halo_set_tuning(0, t0); lib.halo_set_tuning(1, t1)
    best = None
    for it in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        check(lib.halo_msm_dev(0, ctypes.c_void_p(d_pts.data_ptr()), ctypes.c_size_t(0), ctypes.c_void_p(d_sc.data_ptr()), ctypes.c_size_t(k), out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))))
        e1.record(s); e1.synchronize()
        n = lib.halo_profile_read(ms_buf, nm_buf, 24)
        st = {nm_buf[j].decode(): ms_buf[j] for j in range(n)}
        t = e0.elapsed_time(e1)
        if best is None or t < best[0]: best = (t, st)
    print(f"c={c} occ={t0} S={t1}: {best[0]:.3f} ms  " + " ".join(f"{a}={b:.3f}" for a, b in best[1].items()), flush=True)
