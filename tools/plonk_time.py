"""End-to-end PLONK prove/verify timing on the GPU for SURVEY §8(d) config 4: the synthetic multiplication circuit
(w0 = a, w1 = b, w2 = a*b, q_m = 1, q_o = -1, identity permutation, no public inputs, genesis accumulator).
usage: tools/plonk_time.py [log_n ...]      (HALO_TRACE=1 prints the per-round split of the prover)"""
import ctypes, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import halo_b200 as hb
from halo_b200._lib import lib, check
from halo_b200 import consts

hb.init()
for kv in os.environ.get('HALO_TUNE', '').split(','):          # HALO_TUNE="15=1,13=1": tuning keys
    if '=' in kv:
        lib.halo_set_tuning(int(kv.split('=')[0]), int(kv.split('=')[1]))
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
check(lib.halo_set_stream(ctypes.c_void_p(s.cuda_stream)))
u64p = ctypes.POINTER(ctypes.c_uint64)
CURVE = 0
FIELD = 0


def set_pp(n):
    gs = torch.empty((n, 8), dtype=torch.int64, device="cuda")
    check(lib.halo_gen_points_dev(CURVE, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n), ctypes.c_void_p(gs.data_ptr())))
    sh = np.zeros((2, 8), dtype=np.uint64)
    check(lib.halo_gen_points(CURVE, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
    hb.pp.reset(CURVE)
    hb.pp.set_pp(CURVE, gs.cpu().numpy().view(np.uint64), sh[0], sh[1])


def run(lg_n, dense=False, reps=2, satisfied=False):
    n = 1 << lg_n
    set_pp(n)
    t = time.perf_counter(); polys, w_evals = hb.plonk.synthetic_circuit(CURVE, lg_n, dense_gates=dense, satisfied=satisfied); t_build = time.perf_counter() - t
    acc0 = hb.acc.Accumulator.zero(CURVE, n)
    t = time.perf_counter(); circuit = hb.plonk.PlonkCircuit.from_polys(CURVE, polys, 0); t_setup = time.perf_counter() - t
    pub = hb.plonk.PlonkPublicInputs(np.zeros((0, 4), dtype=np.uint64), acc0)
    wit = hb.plonk.PlonkWitness(polys, w_evals)
    best = 1e9
    for r in range(reps):
        if os.environ.get("HALO_TRACE") and r == reps - 1:
            lib.halo_set_tuning(7, 1)
        t = time.perf_counter(); pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit); best = min(best, time.perf_counter() - t)
        lib.halo_set_tuning(7, 0)
    ok = True
    t = time.perf_counter()
    try:
        pi.verify(circuit, pub)
    except hb.Rejected:
        ok = False
    t_verify = time.perf_counter() - t
    circuit.make_resident(polys)
    best_res = 1e9
    for r in range(reps):
        t = time.perf_counter(); pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit); best_res = min(best_res, time.perf_counter() - t)
    circuit.release()
    print(f"n=2^{lg_n} {('all gates' + ('' if satisfied else ', unsatisfied witness')) if dense else 'mul circuit'}: prove {best * 1e3:.1f} ms ({best_res * 1e3:.1f} ms with the circuit resident)  verify {t_verify * 1e3:.1f} ms "
          f"(accepted={ok})  [circuit build {t_build:.1f} s, selector commits {t_setup * 1e3:.0f} ms]", flush=True)
    return best, t_verify


if __name__ == "__main__":
    lgs = [int(x) for x in sys.argv[1:]] or [12, 16]
    for lg in lgs:
        run(lg)
        run(lg, dense=True, satisfied=True)
        run(lg, dense=True, reps=1)
