"""One PC_DL opening over N GPUs (halo_dist_pcdl_open) against the single-GPU opening in the same process.
usage: tools/dist_open_time.py N [log_n ...]     (HALO_TRACE=1: per-round split on rank 0)"""
import ctypes, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch.multiprocessing as mp


def worker(rank, world, idfile, lgs):
    import torch
    import halo_b200 as hb
    from halo_b200._lib import lib, check
    hb.init(rank)
    hb.dist.init_file(rank, world, idfile)
    u64p = ctypes.POINTER(ctypes.c_uint64)
    curve = 0
    for lg in lgs:
        n = 1 << lg
        gs = torch.empty((n, 8), dtype=torch.int64, device=f"cuda:{rank}")
        check(lib.halo_gen_points_dev(curve, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n), ctypes.c_void_p(gs.data_ptr())))
        sh = np.zeros((2, 8), dtype=np.uint64)
        check(lib.halo_gen_points(curve, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
        hb.pp.reset(curve)
        hb.pp.set_pp(curve, gs.cpu().numpy().view(np.uint64), sh[0], sh[1])
        del gs
        rng = np.random.default_rng(11)
        p = rng.integers(0, 1 << 63, size=(n + 1, 4), dtype=np.uint64); p[:, 3] &= np.uint64((1 << 62) - 1)
        C = hb.pcdl.commit(curve, p[:n], n - 1)
        single = dist = 1e9
        for r in range(3):
            t = time.perf_counter(); p1, v1 = hb.pcdl.open_without_eval(curve, p[:n], C, n - 1, p[n], None); single = min(single, time.perf_counter() - t)
        for r in range(4):
            if os.environ.get("HALO_TRACE") and r == 3 and rank == 0:
                lib.halo_set_tuning(7, 1)
            t = time.perf_counter(); pd, vd = hb.dist.pcdl_open(curve, p[:n], C, n - 1, p[n], None); dist = min(dist, time.perf_counter() - t)
            lib.halo_set_tuning(7, 0)
        assert (vd == v1).all() and (pd.w[24 * lg + 12: 24 * lg + 16] == p1.w[24 * lg + 12: 24 * lg + 16]).all(), "final c differs"
        hb.pcdl.Instance.new(curve, C, n - 1, p[n], vd, pd).check()
        if rank == 0:
            print(f"n=2^{lg}: opening (host coefficients in, proof out) on {world} GPUs {dist * 1e3:.2f} ms; single GPU {single * 1e3:.2f} ms", flush=True)
        hb.pp.reset(curve)
    hb.dist.shutdown()


if __name__ == "__main__":
    world = int(sys.argv[1])
    lgs = [int(x) for x in sys.argv[2:]] or [18, 20]
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(worker, args=(world, os.path.join(d, "id"), lgs), nprocs=world, join=True)
