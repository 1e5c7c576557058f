"""Stage times of the MSM pair of one fold-free IPA tail round: 2 MSMs over the same 2^lg frozen generators, half of
each scalar vector zero (complementary halves).  usage: tools/tail_shape.py [lg]"""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 15
T = 1 << lg
pts = torch.empty((T, 8), dtype=torch.int64, device="cuda")
check(lib.halo_gen_points_dev(0, ctypes.c_uint64(3), ctypes.c_size_t(0), ctypes.c_size_t(T), ctypes.c_void_p(pts.data_ptr())))
sc = torch.randint(-(1 << 63), (1 << 63) - 1, (2, T, 4), dtype=torch.int64, device="cuda")
sc[:, :, 3] &= (1 << 62) - 1
sc[0, : T // 2] = 0
sc[1, T // 2:] = 0
s = torch.cuda.Stream(); torch.cuda.set_stream(s)
check(lib.halo_set_stream(ctypes.c_void_p(s.cuda_stream)))
out = np.empty(24, dtype=np.uint64)
u64p = ctypes.POINTER(ctypes.c_uint64)
f = lambda: check(lib.halo_msm_batch_dev(0, ctypes.c_void_p(pts.data_ptr()), ctypes.c_size_t(0), ctypes.c_void_p(sc.data_ptr()),
                                         ctypes.c_size_t(T), ctypes.c_size_t(2), out.ctypes.data_as(u64p)))
for c in [0] + [int(x) for x in sys.argv[2:]]:
    check(lib.halo_msm_set_window(c))
    f(); f()
    import time
    best = 1e9
    for _ in range(5):
        t = time.perf_counter(); f(); best = min(best, time.perf_counter() - t)
    lib.halo_profile_enable(1); f()
    ms = (ctypes.c_float * 16)(); names = (ctypes.c_char_p * 16)()
    ns = lib.halo_profile_read(ms, names, 16)
    lib.halo_profile_enable(0)
    print(f"T=2^{lg} window={c or 'default'}: call {best * 1e3:.3f} ms | " + "  ".join(f"{names[i].decode()} {ms[i]:.3f}" for i in range(ns)), flush=True)
check(lib.halo_msm_set_window(0))
