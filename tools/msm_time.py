"""Quick MSM timing sweep on the GPU (inputs resident in HBM)."""
import ctypes, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
if os.environ.get('HALO_TUNE8'):
    lib.halo_set_tuning(8, int(os.environ['HALO_TUNE8']))
for kv in os.environ.get('HALO_TUNE', '').split(','):          # HALO_TUNE="11=4,8=0": tuning keys
    if '=' in kv:
        lib.halo_set_tuning(int(kv.split('=')[0]), int(kv.split('=')[1]))
lgs = [int(x) for x in sys.argv[1:]] or [12, 16, 18, 20]
kmax = 1 << max(lgs)
d_pts = torch.empty((kmax, 8), dtype=torch.int64, device="cuda")
check(lib.halo_gen_points_dev(0, ctypes.c_uint64(3), ctypes.c_size_t(0), ctypes.c_size_t(kmax), ctypes.c_void_p(d_pts.data_ptr())))
d_sc = torch.randint(-(1 << 63), (1 << 63) - 1, (kmax, 4), dtype=torch.int64, device="cuda")
d_sc[:, 3] &= (1 << 62) - 1
s = torch.cuda.Stream()          # a real stream: handle 0 would mean 'library stream' to halo_set_stream
torch.cuda.set_stream(s)
check(lib.halo_set_stream(ctypes.c_void_p(s.cuda_stream)))
out = np.empty(12, dtype=np.uint64)
for lg in lgs:
    k = 1 << lg
    for c in ([0] if len(sys.argv) > 1 and False else [0]):
        ts = []
        for it in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            check(lib.halo_msm_dev(0, ctypes.c_void_p(d_pts.data_ptr()), ctypes.c_size_t(0), ctypes.c_void_p(d_sc.data_ptr()), ctypes.c_size_t(k), out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))))
            e1.record(s); e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        print(f"k=2^{lg} ms: " + " ".join(f"{x:.3f}" for x in ts) + f"  -> {k / (min(ts) * 1e-3):.3e} pts/s", flush=True)
    if os.environ.get("HALO_STAGES"):
        lib.halo_profile_enable(1)
        for nb in (1, 2):
            outb = np.empty(12 * nb, dtype=np.uint64)
            for it in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(s)
                check(lib.halo_msm_batch_dev(0, ctypes.c_void_p(d_pts.data_ptr()), ctypes.c_size_t(0), ctypes.c_void_p(d_sc.data_ptr()), ctypes.c_size_t(k // nb), ctypes.c_size_t(nb), outb.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))))
                e1.record(s); e1.synchronize()
            ms = (ctypes.c_float * 16)(); names = (ctypes.c_char_p * 16)()
            ns = lib.halo_profile_read(ms, names, 16)
            print(f"   batch={nb} x 2^{lg}/{nb}: total {e0.elapsed_time(e1):.3f} ms  " + "  ".join(f"{names[i].decode()} {ms[i]:.3f}" for i in range(ns)), flush=True)
        lib.halo_profile_enable(0)
