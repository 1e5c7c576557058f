"""AS_DL accumulation over m non-hiding PC_DL instances (Instance::rand_without_hiding shape,
crates/accumulation/src/pcdl.rs:101-108) at fixed n: time acc::prover, acc::verifier, acc::decider
(crates/accumulation/benches/acc.rs:32-83 shapes).  usage: tools/acc_time.py [log_n] [m ...]"""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import halo_b200 as hb
from halo_b200._lib import lib, check

hb.init()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ms_list = [int(x) for x in sys.argv[2:]] or [64, 128, 256, 512, 1024]
n = 1 << lg
u64p = ctypes.POINTER(ctypes.c_uint64)
gs = np.zeros((n, 8), dtype=np.uint64)
check(lib.halo_gen_points(0, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n), gs.ctypes.data_as(u64p)))
sh = np.zeros((2, 8), dtype=np.uint64)
check(lib.halo_gen_points(0, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
hb.pp.set_pp(0, gs, sh[0], sh[1])
rng = np.random.default_rng(1)


def rand_elems(k):
    t = rng.integers(0, 1 << 64, size=(k, 4), dtype=np.uint64)
    t[:, 3] &= np.uint64((1 << 62) - 1)
    return t


mmax = max(ms_list)
t = time.perf_counter()
# distinct instances are opened for the first 64; the rest reuse them round-robin (an opening is 18 ms at 2^16, and
# the accumulation cost does not depend on the instances being distinct)
base = [hb.pcdl.Instance.open(0, rand_elems(n), n - 1, rand_elems(1)[0]) for _ in range(min(64, mmax))]
print(f"n=2^{lg}: opened {len(base)} instances in {time.perf_counter() - t:.1f} s", flush=True)
for m in ms_list:
    qs = [base[i % len(base)] for i in range(m)]
    best = {}
    for rep in range(2):
        if os.environ.get("HALO_TRACE") and rep == 1:
            lib.halo_set_tuning(7, 1)
        t = time.perf_counter(); acc = hb.acc.prover(0, qs); dt = time.perf_counter() - t
        lib.halo_set_tuning(7, 0)
        best["prover"] = min(best.get("prover", 1e9), dt)
        t = time.perf_counter(); hb.acc.verifier(0, qs, acc); dt = time.perf_counter() - t
        best["verifier"] = min(best.get("verifier", 1e9), dt)
        t = time.perf_counter(); hb.acc.decider(0, acc); dt = time.perf_counter() - t
        best["decider"] = min(best.get("decider", 1e9), dt)
    print(f"  m={m:5d}: prover {best['prover'] * 1e3:8.1f} ms  verifier {best['verifier'] * 1e3:8.1f} ms  "
          f"decider {best['decider'] * 1e3:6.1f} ms   ({best['prover'] / m * 1e3:.2f} ms per instance)", flush=True)
