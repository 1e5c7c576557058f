"""crates/plonk/src/plonk/protocol.rs — PlonkProof::naive_prover / verify_succinct / verify over the GPU library.

The circuit, public inputs and witness travel as the flat arrays Trace::consume produces (crates/plonk/src/circuit/
trace.rs:234-283): coefficient columns in the order qs[10], ws[16], rs[15], ids[8], sigmas[8]."""
import ctypes

import numpy as np

from . import _lib, pcdl, acc as acc_mod
from ._lib import lib, check, ptr, as_u64

Q_POLYS, W_POLYS, R_POLYS, S_POLYS, T_POLYS = 10, 16, 15, 8, 16
N_COLUMNS = Q_POLYS + W_POLYS + R_POLYS + 2 * S_POLYS

lib.halo_plonk_proof_words.restype = ctypes.c_size_t


def proof_words(lg_n):
    return int(lib.halo_plonk_proof_words(ctypes.c_uint(lg_n)))


class PlonkCircuit:
    """≙ PlonkCircuit { rows, public_input_count, omega, Cs } (trace.rs:43-50); only Cs.qs is read by the verifier."""

    def __init__(self, curve, rows, public_input_count, C_qs):
        self.curve, self.rows, self.public_input_count = curve, rows, public_input_count
        self.C_qs = np.ascontiguousarray(C_qs, dtype=np.uint64).reshape(Q_POLYS, 12)

    @classmethod
    def from_polys(cls, curve, polys, public_input_count):
        """The selector commitments as Trace::new makes them (trace.rs:188): one batched MSM."""
        polys = np.ascontiguousarray(polys, dtype=np.uint64)
        n = polys.shape[1]
        C_qs = np.stack([pcdl.commit(curve, polys[i], n - 1) for i in range(Q_POLYS)])
        return cls(curve, n, public_input_count, C_qs)


class PlonkPublicInputs:
    """≙ PlonkPublicInputs { public_inputs, acc_prev } (trace.rs:37-41)."""

    def __init__(self, public_inputs, acc_prev):
        self.public_inputs = as_u64(public_inputs, 4)
        self.acc_prev = acc_prev


class PlonkWitness:
    """≙ PlonkWitness { omega, polys, w_evals } (trace.rs:52-66)."""

    def __init__(self, polys, w_evals):
        self.polys = np.ascontiguousarray(polys, dtype=np.uint64)
        assert self.polys.shape[0] == N_COLUMNS and self.polys.shape[2] == 4
        self.w_evals = np.ascontiguousarray(w_evals, dtype=np.uint64)[:3]


class PlonkProof:
    """≙ PlonkProof { vs, Cs, pis, acc_next } (protocol.rs:29-61), kept as the flat word array of the C ABI."""

    def __init__(self, curve, lg_n, words):
        self.curve, self.lg_n = curve, lg_n
        self.w = np.ascontiguousarray(words, dtype=np.uint64)
        assert self.w.shape[0] == proof_words(lg_n)

    # ---- views ----
    def evals(self):
        v = self.w[:78 * 4].reshape(78, 4)
        return dict(ws=v[0:16], rs=v[16:31], qs=v[31:41], ts=v[41:57], ids=v[57:65], sigmas=v[65:73], z=v[73],
                    z_omega=v[74], w_omegas=v[75:78])

    def commitments(self):
        c = self.w[78 * 4:78 * 4 + 33 * 12].reshape(33, 12)
        return dict(ws=c[0:16], ts=c[16:32], z=c[32])

    @property
    def acc_next(self):
        return acc_mod.Accumulator(pcdl.Instance(self.curve, self.w[-int(lib.halo_instance_words(self.lg_n)):]))

    # ---- protocol.rs:64-355 ----
    @classmethod
    def naive_prover(cls, circuit, public_inputs, witness):
        _lib.ensure_init()
        n = circuit.rows
        lg_n = n.bit_length() - 1
        if n != 1 << lg_n or witness.polys.shape[1] != n:
            raise _lib.PreconditionError(_lib.HALO_EINVAL, "rows must be a power of two matching the columns")
        out = np.zeros(proof_words(lg_n), dtype=np.uint64)
        pub = public_inputs.public_inputs
        check(lib.halo_plonk_prove(circuit.curve, ctypes.c_uint(lg_n), ptr(witness.polys.reshape(-1)),
                                   ptr(witness.w_evals.reshape(-1)), ptr(pub.reshape(-1)) if pub.shape[0] else None,
                                   ctypes.c_size_t(pub.shape[0]), ptr(public_inputs.acc_prev.q.w), ptr(out)))
        return cls(circuit.curve, lg_n, out)

    # ---- protocol.rs:357-502 ----
    def _verify(self, circuit, public_inputs, succinct_only):
        _lib.ensure_init()
        pub = public_inputs.public_inputs
        check(lib.halo_plonk_verify(circuit.curve, ctypes.c_uint(self.lg_n), ptr(circuit.C_qs.reshape(-1)),
                                    ptr(pub.reshape(-1)) if pub.shape[0] else None, ctypes.c_size_t(pub.shape[0]),
                                    ctypes.c_size_t(circuit.public_input_count), ptr(public_inputs.acc_prev.q.w),
                                    ptr(self.w), ctypes.c_int(1 if succinct_only else 0)))

    def verify_succinct(self, circuit, public_inputs):
        """Raises Rejected where the reference's ensure! fails."""
        self._verify(circuit, public_inputs, True)

    def verify(self, circuit, public_inputs):
        self._verify(circuit, public_inputs, False)
