"""SURVEY §8(d) configs 2 and 3 as one JSON document (inputs resident in HBM, CUDA events, best of 5 after a warm-up):
  config 2: Pallas and Vesta variable-base MSM, k = 2^12 .. 2^22, random points and uniform scalars
  config 3: Fp / Fq NTT, iNTT and coset NTT, N = 2^12 .. 2^22, batches of 1, 16 and 42 polynomials
Every line carries the fraction of the live modmul peak (halo_bench_modmul), the roof of this path.
usage: tools/sweep.py [out.json]"""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200
from halo_b200._lib import lib, check

halo_b200.init()
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
check(lib.halo_set_stream(ctypes.c_void_p(s.cuda_stream)))
u64p = ctypes.POINTER(ctypes.c_uint64)


def ev_time(fn, reps=5):
    fn()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s); fn(); e1.record(s); e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def rand_elems(n, seed):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    t = torch.randint(-(1 << 63), (1 << 63) - 1, (n, 4), dtype=torch.int64, device="cuda", generator=g)
    t[:, 3] &= (1 << 62) - 1
    return t


def modmul_peak(field):
    ms = ctypes.c_float(0); out = (ctypes.c_uint64 * 4)()
    blocks, threads, iters = 148 * 16, 128, 2000
    check(lib.halo_bench_modmul(field, blocks, threads, iters, ctypes.byref(ms), out))
    return blocks * threads * iters * 4 / (ms.value * 1e-3)


doc = {"modmul_peak_per_s": {"Fp": modmul_peak(0), "Fq": modmul_peak(1)}, "msm": [], "ntt": []}
LG_MAX = 22

# ---- config 2 ----
for curve, name in ((0, "pallas"), (1, "vesta")):
    kmax = 1 << LG_MAX
    pts = torch.empty((kmax, 8), dtype=torch.int64, device="cuda")
    check(lib.halo_gen_points_dev(curve, ctypes.c_uint64(3), ctypes.c_size_t(0), ctypes.c_size_t(kmax), ctypes.c_void_p(pts.data_ptr())))
    sc = rand_elems(kmax, 11 + curve)
    out = np.empty(12, dtype=np.uint64)
    peak = doc["modmul_peak_per_s"]["Fq" if curve == 0 else "Fp"]          # base field of the curve
    for lg in range(12, LG_MAX + 1):
        k = 1 << lg
        f = lambda: check(lib.halo_msm_dev(curve, ctypes.c_void_p(pts.data_ptr()), ctypes.c_size_t(0), ctypes.c_void_p(sc.data_ptr()),
                                           ctypes.c_size_t(k), out.ctypes.data_as(u64p)))
        ms = ev_time(f)
        row = {"curve": name, "log_k": lg, "ms": round(ms, 4), "points_per_s": k / (ms * 1e-3)}
        doc["msm"].append(row)
        print(json.dumps(row), flush=True)
    del pts, sc

# ---- config 3 ----
for field, name in ((0, "Fp"), (1, "Fq")):
    peak = doc["modmul_peak_per_s"][name]
    coset = np.zeros(4, dtype=np.uint64)
    five = halo_b200.group.field_op(field, "from_canonical", np.array([[5, 0, 0, 0]], dtype=np.uint64))[0]
    for lg in range(12, LG_MAX + 1, 2):
        for b in (1, 16, 42):
            if (b << lg) > (1 << 27):
                continue
            x = rand_elems(b << lg, lg + b)
            for kind, inv, cs in (("ntt", 0, None), ("intt", 1, None), ("coset_ntt", 0, five)):
                if field == 1 and kind != "ntt":
                    continue
                f = lambda: check(lib.halo_ntt_dev(field, ctypes.c_void_p(x.data_ptr()), ctypes.c_uint(lg), ctypes.c_size_t(b), inv,
                                                   cs.ctypes.data_as(u64p) if cs is not None else None))
                ms = ev_time(f)
                n = b << lg
                bf = n * lg / 2
                row = {"field": name, "kind": kind, "log_n": lg, "batch": b, "ms": round(ms, 4), "elements_per_s": n / (ms * 1e-3),
                       "gbutterfly_per_s": bf / (ms * 1e-3) / 1e9, "frac_of_modmul_peak": bf / (ms * 1e-3) / peak,
                       "hbm_gb_per_s_one_pass_equivalent": 64 * n / (ms * 1e-3) / 1e9}
                doc["ntt"].append(row)
                print(json.dumps(row), flush=True)
            del x

if len(sys.argv) > 1:
    json.dump(doc, open(sys.argv[1], "w"), indent=1)
