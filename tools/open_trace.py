"""Per-round phase trace of one IPA opening (tuning knob 7)."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n = 1 << lg
u64p = ctypes.POINTER(ctypes.c_uint64)
gs = torch.empty((n, 8), dtype=torch.int64, device="cuda")
check(lib.halo_gen_points_dev(0, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n), ctypes.c_void_p(gs.data_ptr())))
sh = np.zeros((2, 8), dtype=np.uint64)
check(lib.halo_gen_points(0, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
halo_b200.pp.set_pp(0, gs.cpu().numpy().view(np.uint64), sh[0], sh[1])
t = torch.randint(-(1 << 63), (1 << 63) - 1, (n + 1, 4), dtype=torch.int64); t[:, 3] &= (1 << 62) - 1
p = t.numpy().view(np.uint64)
q = halo_b200.pcdl.Instance.open(0, p[:n], n - 1, p[n])
lib.halo_set_tuning(7, 1)
q = halo_b200.pcdl.Instance.open(0, p[:n], n - 1, p[n])
