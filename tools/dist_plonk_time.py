"""ONE PLONK proof over N GPUs (halo_dist_plonk_prove): timing against the single-GPU prover of the same process.
usage: tools/dist_plonk_time.py N [log_n ...]     (HALO_TRACE=1: per-round split on rank 0)"""
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
        for dense in (False, True):
            polys, w_evals = hb.plonk.synthetic_circuit(curve, lg, dense_gates=dense, satisfied=dense)
            acc0 = hb.acc.Accumulator.zero(curve, n)
            circuit = hb.plonk.PlonkCircuit.from_polys(curve, polys, 0)
            pub = hb.plonk.PlonkPublicInputs(np.zeros((0, 4), dtype=np.uint64), acc0)
            wit = hb.plonk.PlonkWitness(polys, w_evals)
            single = 1e9
            for r in range(2):
                t = time.perf_counter(); p1 = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit); single = min(single, time.perf_counter() - t)
            res = {}
            for resident in (False, True):
                if resident:
                    circuit.make_resident(polys)
                best = 1e9
                for r in range(3):
                    if os.environ.get("HALO_TRACE") and r == 2 and rank == 0:
                        lib.halo_set_tuning(7, 1)
                    t = time.perf_counter()
                    pd = hb.dist.plonk_prove(circuit, pub, wit) if rank == 0 else hb.dist.plonk_prove(circuit, 0, None)
                    best = min(best, time.perf_counter() - t)
                    lib.halo_set_tuning(7, 0)
                if resident:
                    circuit.release()
                res[resident] = best
                assert (pd.w[:78 * 4] == p1.w[:78 * 4]).all(), "evaluations differ from the single-GPU proof"
                if os.environ.get("HALO_CHECK"):
                    L = lg
                    ew = 24 * L + 12 + 4 + 1 + 12 + 4
                    o = 78 * 4 + 33 * 12
                    parts = {"Cs": (78 * 4, o), "pi_r": (o, o + ew), "pi_ro": (o + ew, o + 2 * ew), "acc": (o + 2 * ew, len(pd.w))}
                    for nm, (a, b) in parts.items():
                        same = bool((pd.w[a:b] == p1.w[a:b]).all())
                        print(f"rank {rank} lg {lg} dense {dense} resident {resident}: {nm} words equal to single-GPU: {same}", flush=True)
                    if not dense:
                        try:
                            pd.verify(circuit, pub); print(f"rank {rank}: dist proof verifies", flush=True)
                        except Exception as e:
                            print(f"rank {rank}: dist proof REJECTED {e}", flush=True)
                        try:
                            p1.verify(circuit, pub); print(f"rank {rank}: single proof verifies", flush=True)
                        except Exception as e:
                            print(f"rank {rank}: single proof REJECTED {e}", flush=True)
            if rank == 0:
                print(f"n=2^{lg} {'all gates' if dense else 'mul circuit'}: one proof on {world} GPUs {res[False] * 1e3:.1f} ms "
                      f"({res[True] * 1e3:.1f} ms circuit-resident); single GPU in this process {single * 1e3:.1f} ms", flush=True)
        hb.pp.reset(curve)
    hb.dist.shutdown()


if __name__ == "__main__":
    world = int(sys.argv[1])
    lgs = [int(x) for x in sys.argv[2:]] or [16, 20]
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(worker, args=(world, os.path.join(d, "id"), lgs), nprocs=world, join=True)
