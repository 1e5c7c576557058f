"""A few device-resident batched NTTs (for ncu)."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
b = int(sys.argv[2]) if len(sys.argv) > 2 else 16
x = torch.randint(-(1 << 63), (1 << 63) - 1, ((b << lg), 4), dtype=torch.int64, device="cuda")
x[:, 3] &= (1 << 62) - 1
for it in range(3):
    check(lib.halo_ntt_dev(0, ctypes.c_void_p(x.data_ptr()), ctypes.c_uint(lg), ctypes.c_size_t(b), it & 1, None))
check(lib.halo_device_synchronize())
print("ok")
