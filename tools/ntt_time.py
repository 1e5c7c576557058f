"""Device-resident batched NTT timings (CUDA events, best of 5).  usage: tools/ntt_time.py [lg ...]"""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
s = torch.cuda.Stream(); torch.cuda.set_stream(s)
check(lib.halo_set_stream(ctypes.c_void_p(s.cuda_stream)))
for lg in [int(a) for a in sys.argv[1:]] or [16, 18, 20, 22]:
    for b in (1, 16, 42):
        if (b << lg) > (1 << 27):
            continue
        x = torch.randint(-(1 << 63), (1 << 63) - 1, ((b << lg), 4), dtype=torch.int64, device="cuda")
        x[:, 3] &= (1 << 62) - 1
        res = []
        for inv in (0, 1):
            ts = []
            for it in range(6):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(s)
                check(lib.halo_ntt_dev(0, ctypes.c_void_p(x.data_ptr()), ctypes.c_uint(lg), ctypes.c_size_t(b), inv, None))
                e1.record(s); e1.synchronize()
                ts.append(e0.elapsed_time(e1))
            res.append(min(ts[1:]))
        bf = b * (1 << lg) * lg / 2
        print(f"2^{lg} x {b:2d}: forward {res[0]:8.3f} ms ({bf / res[0] / 1e6:6.1f} Gbutterfly/s)   inverse {res[1]:8.3f} ms", flush=True)
        del x
