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
        self.handle = 0

    @classmethod
    def from_polys(cls, curve, polys, public_input_count, resident=False):
        """The selector commitments as Trace::new makes them (trace.rs:188): one batched MSM.
        resident=True also keeps the circuit's own columns (qs, rs, ids, sigmas) on the device, so that every proof
        of this circuit uploads only its witness columns."""
        polys = np.ascontiguousarray(polys, dtype=np.uint64)
        n = polys.shape[1]
        C_qs = np.stack([pcdl.commit(curve, polys[i], n - 1) for i in range(Q_POLYS)])
        c = cls(curve, n, public_input_count, C_qs)
        if resident:
            c.make_resident(polys)
        return c

    def make_resident(self, polys):
        """halo_plonk_circuit_load: the 41 circuit columns of `polys` (57, n, 4) stay in HBM until release()."""
        _lib.ensure_init()
        polys = np.ascontiguousarray(polys, dtype=np.uint64)
        assert polys.shape == (N_COLUMNS, self.rows, 4)
        self.release()
        h = ctypes.c_int(0)
        check(lib.halo_plonk_circuit_load(self.curve, ctypes.c_uint(self.rows.bit_length() - 1), ptr(polys.reshape(-1)),
                                          ctypes.byref(h)))
        self.handle = h.value

    def release(self):
        if getattr(self, "handle", 0):
            lib.halo_plonk_circuit_free(ctypes.c_int(self.handle))
            self.handle = 0

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


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
        if circuit.handle:                                  # circuit columns resident: only the witness crosses the link
            w_polys = witness.polys[Q_POLYS:Q_POLYS + W_POLYS]
            check(lib.halo_plonk_prove_circuit(ctypes.c_int(circuit.handle), ptr(w_polys.reshape(-1)),
                                               ptr(witness.w_evals.reshape(-1)), ptr(pub.reshape(-1)) if pub.shape[0] else None,
                                               ctypes.c_size_t(pub.shape[0]), ptr(public_inputs.acc_prev.q.w), ptr(out)))
        else:
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


def gate_residual(curve, polys, w_evals, public_inputs=None):
    """f_GC on H (protocol.rs:170-241 without the permutation terms): (n, 4) values f_GC(omega^j); all zero exactly when
    every gate row is satisfied."""
    _lib.ensure_init()
    n = polys.shape[1]
    lg_n = n.bit_length() - 1
    pub = public_inputs if public_inputs is not None else np.zeros((0, 4), dtype=np.uint64)
    out = np.zeros((n, 4), dtype=np.uint64)
    check(lib.halo_plonk_gate_residual(curve, ctypes.c_uint(lg_n), ptr(np.ascontiguousarray(polys).reshape(-1)),
                                       ptr(np.ascontiguousarray(w_evals).reshape(-1)), ptr(pub.reshape(-1)) if pub.shape[0] else None,
                                       ctypes.c_size_t(pub.shape[0]), ptr(out.reshape(-1))))
    return out


def synthetic_circuit(curve, lg_n, seed=1, dense_gates=False, satisfied=False):
    """The synthetic workload of SURVEY §8(d) config 4, built with GPU arithmetic: rows w0 = a, w1 = b, w2 = a*b with
    q_m = 1, q_o = -1, identity permutation, no public inputs.  dense_gates=True additionally fills the five custom
    selectors, the round-constant columns and the remaining witnesses with random values (every gate expression
    is then evaluated; the circuit is no longer satisfied).  satisfied=True (with dense_gates) then solves the constant
    selector row by row, q_c := -(everything else of f_GC), so that the all-gates circuit IS satisfied -- what an honest
    prover's circuit is -- and the quotient takes the 8-coset path.  Returns (polys (57, n, 4), w_evals (3, n, 4))."""
    from . import consts, group, poly
    field = curve
    n = 1 << lg_n
    rng = np.random.default_rng(seed)

    def rand_elems(k):
        t = rng.integers(0, 1 << 64, size=(k, 4), dtype=np.uint64)
        t[:, 3] &= np.uint64((1 << 62) - 1)            # < 2^254 < p: a valid Montgomery residue
        return t

    one = consts.one(field)
    neg_one = group.field_op(field, "sub", np.zeros((1, 4), dtype=np.uint64), one.reshape(1, 4))[0]
    cols = np.zeros((N_COLUMNS, n, 4), dtype=np.uint64)
    a, b = rand_elems(n), rand_elems(n)
    cols[10], cols[11] = a, b
    cols[12] = group.field_op(field, "mul", a, b)
    cols[3] = one                                       # q_m
    cols[2] = neg_one                                   # q_o
    if dense_gates:
        for i in list(range(5, 10)) + list(range(13, 26)) + list(range(26, 41)):
            cols[i] = rand_elems(n)
    idx = np.zeros((S_POLYS * n, 4), dtype=np.uint64)
    idx[:, 0] = np.arange(1, S_POLYS * n + 1, dtype=np.uint64)    # SlotId::to_scalar: row + (column-1)*n, 1-based
    ids = group.field_op(field, "from_canonical", idx)
    cols[41:49] = ids.reshape(S_POLYS, n, 4)
    cols[49:57] = cols[41:49]
    evals = np.roll(cols, 1, axis=1)                    # Evals::from_vec_and_domain (poly.rs:21-24)
    polys = np.empty_like(cols)
    for lo in range(0, N_COLUMNS, 19):
        polys[lo:lo + 19] = poly.ntt(field, evals[lo:lo + 19], inverse=True)
    if dense_gates and satisfied:
        res = gate_residual(curve, polys, evals[10:13].copy())
        qc = group.field_op(field, "sub", evals[4], res)                  # f_GC is linear in q_c with coefficient one
        polys[4] = poly.ntt(field, qc.reshape(1, n, 4), inverse=True)[0]
    return polys, evals[10:13].copy()


class Trace:
    """≙ Trace::new / Trace::consume (crates/plonk/src/circuit/trace.rs:137-283): column vectors -> Evals (row i at
    omega^(i+1)) -> coefficient polynomials by ONE batched iNTT of the 57 columns, and the circuit commitments
    (10 + 15 + 8 + 8) by batched MSMs over the resident generators instead of 41 sequential commits.

    copy_constraints: list of equivalence classes of 1-based (row, column) slots, as build_sigma takes them
    (trace.rs:64-104); ws / rs / qs: (16|15|10, n, 4) Montgomery column vectors; public_inputs: (k, 4)."""

    def __init__(self, curve, copy_constraints, public_inputs, ws, rs, qs, n, acc_prev, circuit=None):
        from . import group, poly, pedersen
        _lib.ensure_init()
        field = curve
        lg_n = n.bit_length() - 1
        if n != 1 << lg_n:
            raise _lib.PreconditionError(_lib.HALO_EINVAL, "rows must be a power of two (build_sigma asserts it)")
        ws, rs, qs = (np.ascontiguousarray(x, dtype=np.uint64) for x in (ws, rs, qs))
        assert ws.shape == (W_POLYS, n, 4) and rs.shape == (R_POLYS, n, 4) and qs.shape == (Q_POLYS, n, 4)
        # build_sigma: identity permutation over the 8 n slots, then one cycle per equivalence class
        sigma = np.arange(S_POLYS * n, dtype=np.uint64)
        slot = lambda rc: (rc[0] - 1) + (rc[1] - 1) * n                       # SlotId::to_usize
        for wires in copy_constraints:
            if len(wires) <= 1:
                continue
            for i, frm in enumerate(wires):
                sigma[slot(frm)] = slot(wires[(i + 1) % len(wires)])
        self.sigma = sigma
        canon = np.zeros((2 * S_POLYS * n, 4), dtype=np.uint64)
        canon[:S_POLYS * n, 0] = np.arange(1, S_POLYS * n + 1, dtype=np.uint64)   # SlotId::to_scalar, 1-based
        canon[S_POLYS * n:, 0] = sigma + np.uint64(1)
        idsig = group.field_op(field, "from_canonical", canon).reshape(2 * S_POLYS, n, 4)
        cols = np.concatenate([qs, ws, rs, idsig])                            # flat order of the C ABI
        evals = np.roll(cols, 1, axis=1)                                      # Evals::from_vec_and_domain
        polys = np.empty_like(cols)
        for lo in range(0, N_COLUMNS, 19):
            polys[lo:lo + 19] = poly.ntt(field, evals[lo:lo + 19], inverse=True)
        self.curve, self.rows = curve, n
        self.polys, self.w_evals = polys, evals[Q_POLYS:Q_POLYS + W_POLYS].copy()
        self.public_inputs = as_u64(public_inputs, 4)
        self.acc_prev = acc_prev
        if circuit is not None:
            self.C_qs, self.C_rs, self.C_ids, self.C_sigmas = circuit.C_qs, circuit.C_rs, circuit.C_ids, circuit.C_sigmas
        else:
            if pp_len(curve) < n:
                raise _lib.ParamsError(_lib.HALO_EPARAMS, "public parameters not loaded for this many rows")
            C = group.msm_batch_resident(curve, np.concatenate([polys[:Q_POLYS], polys[26:41], polys[41:57]]))
            self.C_qs, self.C_rs, self.C_ids, self.C_sigmas = C[:10], C[10:25], C[25:33], C[33:41]
        # C_public_inputs = pedersen::commit(None, &pp.Gs, &public_inputs) (trace.rs:195-196)
        k = self.public_inputs.shape[0]
        self.C_public_inputs = pedersen.commit(curve, None, (0, pp_len(curve)), self.public_inputs) if k else None

    def consume(self):
        circuit = PlonkCircuit(self.curve, self.rows, self.public_inputs.shape[0], self.C_qs)
        circuit.C_rs, circuit.C_ids, circuit.C_sigmas = self.C_rs, self.C_ids, self.C_sigmas
        return circuit, PlonkPublicInputs(self.public_inputs, self.acc_prev), PlonkWitness(self.polys, self.w_evals)


def pp_len(curve):
    lib.halo_pp_len.restype = ctypes.c_size_t
    return int(lib.halo_pp_len(curve))
