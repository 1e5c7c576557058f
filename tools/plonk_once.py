"""One PLONK proof of the synthetic multiplication circuit (all custom gates switched on with HALO_DENSE=1), for
profiler runs.  usage: tools/plonk_once.py [log_n]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import halo_b200 as hb
from halo_b200._lib import lib, check

hb.init()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 16
n = 1 << lg
u64p = ctypes.POINTER(ctypes.c_uint64)
gs = np.zeros((n, 8), dtype=np.uint64)
check(lib.halo_gen_points(0, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n), gs.ctypes.data_as(u64p)))
sh = np.zeros((2, 8), dtype=np.uint64)
check(lib.halo_gen_points(0, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
hb.pp.set_pp(0, gs, sh[0], sh[1])
polys, w_evals = hb.plonk.synthetic_circuit(0, lg, dense_gates=bool(os.environ.get("HALO_DENSE")))
acc0 = hb.acc.Accumulator.zero(0, n)
circuit = hb.plonk.PlonkCircuit.from_polys(0, polys, 0)
pub = hb.plonk.PlonkPublicInputs(np.zeros((0, 4), dtype=np.uint64), acc0)
pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, hb.plonk.PlonkWitness(polys, w_evals))
print("proved", lg, "words", pi.w.shape[0])
