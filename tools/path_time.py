"""Timing of the other rows of the hot path on the GPU (no oracle): NTT sweep, IPA opening, accumulation decider.
usage: tools/path_time.py [max_log_n]"""
import ctypes, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200
from halo_b200._lib import lib, check
halo_b200.init()
s = torch.cuda.Stream()          # a real stream: handle 0 would mean 'library stream' to halo_set_stream
torch.cuda.set_stream(s)
check(lib.halo_set_stream(ctypes.c_void_p(s.cuda_stream)))
MAXLG = int(sys.argv[1]) if len(sys.argv) > 1 else 20
u64p = ctypes.POINTER(ctypes.c_uint64)


def rand_scalars(n, seed):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    t = torch.randint(-(1 << 63), (1 << 63) - 1, (n, 4), dtype=torch.int64, device="cuda", generator=g)
    t[:, 3] &= (1 << 62) - 1
    return t


def ev_time(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s); fn(); e1.record(s); e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


print("== NTT (device-resident, in place), field Fp")
for lg in (12, 16, 20, 22, 24):
    if lg > MAXLG + 4: break
    for b in (1, 16):
        if (b << lg) > (1 << 26): continue
        x = rand_scalars(b << lg, lg)
        f = lambda: check(lib.halo_ntt_dev(0, ctypes.c_void_p(x.data_ptr()), ctypes.c_uint(lg), ctypes.c_size_t(b), 0, None))
        f(); ms = ev_time(f)
        n = b << lg
        print(f"  N=2^{lg} batch={b}: {ms:.3f} ms  {n / ms / 1e3:.1f} Melem/s  {n * lg / 2 / ms / 1e6:.2f} Gbutterfly/s  hbm {64 * n * 1e-6 / ms:.0f} GB/s/pass-equivalent", flush=True)

print("== IPA opening / check (Pallas), synthetic generators")
for lg in (12, 16, MAXLG):
    n = 1 << lg
    gs = torch.empty((n, 8), dtype=torch.int64, device="cuda")
    check(lib.halo_gen_points_dev(0, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n), ctypes.c_void_p(gs.data_ptr())))
    sh = np.zeros((2, 8), dtype=np.uint64)
    check(lib.halo_gen_points(0, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
    halo_b200.pp.reset(0)
    halo_b200.pp.set_pp(0, gs.cpu().numpy().view(np.uint64), sh[0], sh[1])
    p = rand_scalars(n, 7).cpu().numpy().view(np.uint64)
    z = rand_scalars(1, 8).cpu().numpy().view(np.uint64)[0]
    t = time.perf_counter(); C = halo_b200.pcdl.commit(0, p, n - 1); t_commit = time.perf_counter() - t
    t = time.perf_counter(); q = halo_b200.pcdl.Instance.open(0, p, n - 1, z); t_open = time.perf_counter() - t
    t = time.perf_counter(); q = halo_b200.pcdl.Instance.open(0, p, n - 1, z); t_open = min(t_open, time.perf_counter() - t)
    t = time.perf_counter(); q.succinct_check(); t_sc = time.perf_counter() - t
    t = time.perf_counter(); q.check(); t_check = time.perf_counter() - t
    t = time.perf_counter(); acc = halo_b200.acc.prover(0, [q, q, q]); t_acc = time.perf_counter() - t
    t = time.perf_counter(); halo_b200.acc.decider(0, acc); t_dec = time.perf_counter() - t
    print(f"  n=2^{lg}: commit(e2e) {t_commit*1e3:.1f} ms  Instance.open {t_open*1e3:.1f} ms  succinct_check {t_sc*1e3:.1f} ms  "
          f"check {t_check*1e3:.1f} ms  acc.prover(m=3) {t_acc*1e3:.1f} ms  decider {t_dec*1e3:.1f} ms", flush=True)
