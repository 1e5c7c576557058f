"""Multi-GPU calls of the C ABI (include/halo_b200.h, halo_dist_*): one process per GPU, the NCCL communicator lives
INSIDE the library, so a host without torch.distributed (the reference's Rust callers) gets the same sharding.

Every function here is collective: all ranks call it in the same order and receive the same result.
  msm / msm_global   ≙ point_dot_affine (crates/group/src/group.rs:48-50) sharded by point range
  acc_prover / acc_verifier / acc_decider  ≙ crates/accumulation/src/acc.rs:178-217 sharded by instance
  ntt                ≙ batched Evals::from_poly / interpolate (crates/group/src/poly.rs:56-64,133-139) by polynomial
  pcdl_open          ≙ pcdl::open_without_eval (crates/accumulation/src/pcdl.rs:326-473), upper rounds on interleaved slices
  plonk_prove        ≙ PlonkProof::naive_prover (crates/plonk/src/plonk/protocol.rs:64-355): ONE proof over all GPUs
The 128-byte NCCL id travels over whatever channel the launcher has; `init_torch` uses a torch.distributed broadcast
(any backend), `init_file` a file on a shared path."""
import ctypes
import os
import time

import numpy as np

from . import _lib, pcdl
from . import acc as _acc
from ._lib import lib, check, ptr, as_u64

_u8p = ctypes.POINTER(ctypes.c_uint8)


def unique_id():
    """Rank 0: a fresh NCCL id (128 bytes)."""
    buf = (ctypes.c_uint8 * 128)()
    check(lib.halo_dist_unique_id(buf))
    return bytes(buf)


def init(rank, world, id_bytes):
    """halo_dist_init: joins the communicator.  halo_init(device) must have been called for this rank's GPU."""
    _lib.ensure_init()
    assert len(id_bytes) == 128
    buf = (ctypes.c_uint8 * 128).from_buffer_copy(id_bytes)
    check(lib.halo_dist_init(ctypes.c_int(rank), ctypes.c_int(world), buf))


def init_torch(tdist):
    """Rendezvous through an already initialised torch.distributed group (gloo or nccl)."""
    import torch
    rank, world = tdist.get_rank(), tdist.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device()) if tdist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        t = torch.frombuffer(bytearray(unique_id()), dtype=torch.uint8).to(dev)
    tdist.broadcast(t, 0)
    init(rank, world, bytes(t.cpu().numpy().tobytes()))


def init_file(rank, world, path, timeout=120.0):
    """Rendezvous through a file: rank 0 writes the id (atomically), the others wait for it."""
    if rank == 0:
        tmp = path + ".tmp"
        with open(tmp, "wb") as f:
            f.write(unique_id())
        os.replace(tmp, path)
    t0 = time.time()
    while not os.path.exists(path):
        if time.time() - t0 > timeout:
            raise TimeoutError(f"no NCCL id at {path}")
        time.sleep(0.01)
    init(rank, world, open(path, "rb").read())


def shutdown():
    check(lib.halo_dist_shutdown())


def info():
    r, w = ctypes.c_int(0), ctypes.c_int(1)
    rc = lib.halo_dist_info(ctypes.byref(r), ctypes.byref(w))
    return (r.value, w.value) if rc == 0 else None


def msm(curve, xs_local, Gs_local=None, base_off=0):
    """This rank's slice of a point-range-sharded MSM; Gs_local=None uses the resident generators Gs[base_off..]."""
    xs = as_u64(xs_local, 4)
    Gs = as_u64(Gs_local, 8) if Gs_local is not None else None
    k = xs.shape[0] if Gs is None else min(xs.shape[0], Gs.shape[0])
    out = np.empty(12, dtype=np.uint64)
    check(lib.halo_dist_msm(curve, ptr(Gs) if Gs is not None and k else None, ctypes.c_size_t(base_off),
                            ptr(xs) if k else None, ctypes.c_size_t(k), ptr(out)))
    return out


def msm_global(curve, xs, Gs=None, base_off=0):
    """≙ point_dot_affine with the WHOLE problem on every rank; the library slices it by rank."""
    xs = as_u64(xs, 4)
    G = as_u64(Gs, 8) if Gs is not None else None
    k = xs.shape[0] if G is None else min(xs.shape[0], G.shape[0])
    out = np.empty(12, dtype=np.uint64)
    check(lib.halo_dist_msm_global(curve, ptr(G) if G is not None and k else None, ctypes.c_size_t(base_off),
                                   ptr(xs) if k else None, ctypes.c_size_t(k), ptr(out)))
    return out


def acc_prover(curve, qs):
    flat = _acc._pack(qs)
    out = np.zeros(flat.shape[1] if qs else 0, dtype=np.uint64)
    check(lib.halo_dist_acc_prover(curve, ptr(flat.reshape(-1)) if qs else None, ctypes.c_size_t(len(qs)), ptr(out)))
    return _acc.Accumulator(pcdl.Instance(curve, out))


def acc_verifier(curve, qs, acc):
    flat = _acc._pack(qs)
    check(lib.halo_dist_acc_verifier(curve, ptr(flat.reshape(-1)) if qs else None, ctypes.c_size_t(len(qs)), ptr(acc.q.w)))


def acc_decider(curve, acc):
    check(lib.halo_dist_acc_decider(curve, ptr(acc.q.w)))


def ntt(field, data, inverse=False, coset=None):
    a = np.ascontiguousarray(data, dtype=np.uint64)
    shape = a.shape
    a = a.reshape(-1, shape[-2], 4).copy()
    n = a.shape[1]
    assert n & (n - 1) == 0, "length must be a power of two"
    g = as_u64(coset, 4).reshape(4) if coset is not None else None
    check(lib.halo_dist_ntt(field, ptr(a.reshape(-1, 4)), ctypes.c_uint(n.bit_length() - 1), ctypes.c_size_t(a.shape[0]),
                            ctypes.c_int(1 if inverse else 0), ptr(g)))
    return a.reshape(shape)


def plonk_prove(circuit, public_inputs=None, witness=None):
    """≙ PlonkProof::naive_prover as ONE proof over the GPUs of the communicator.  `circuit` (shape / resident handle)
    is needed on every rank; `public_inputs` and `witness` are read on rank 0 only (other ranks may pass None, but
    every rank must know how many public inputs there are: pass the same `public_inputs`, or an int count)."""
    from . import plonk as _plonk
    _lib.ensure_init()
    n = circuit.rows
    lg_n = n.bit_length() - 1
    out = np.zeros(_plonk.proof_words(lg_n), dtype=np.uint64)
    if isinstance(public_inputs, int):
        n_pub, pub, acc_prev = public_inputs, None, None
    else:
        pub = public_inputs.public_inputs
        n_pub, acc_prev = pub.shape[0], public_inputs.acc_prev.q.w
    pub_p = ptr(pub.reshape(-1)) if pub is not None and n_pub else None
    acc_p = ptr(acc_prev) if acc_prev is not None else None
    if circuit.handle:
        w_polys = witness.polys[_plonk.Q_POLYS:_plonk.Q_POLYS + _plonk.W_POLYS] if witness is not None else None
        check(lib.halo_dist_plonk_prove_circuit(ctypes.c_int(circuit.handle), ptr(w_polys.reshape(-1)) if w_polys is not None else None,
                                                ptr(witness.w_evals.reshape(-1)) if witness is not None else None, pub_p,
                                                ctypes.c_size_t(n_pub), acc_p, ptr(out)))
    else:
        check(lib.halo_dist_plonk_prove(circuit.curve, ctypes.c_uint(lg_n),
                                        ptr(witness.polys.reshape(-1)) if witness is not None else None,
                                        ptr(witness.w_evals.reshape(-1)) if witness is not None else None, pub_p,
                                        ctypes.c_size_t(n_pub), acc_p, ptr(out)))
    return _plonk.PlonkProof(circuit.curve, lg_n, out)


def pcdl_open(curve, coeffs, C, d, z, v=None, w=None, hide_q=None, hide_wbar=None):
    """≙ pcdl::open_without_eval over the GPUs of the communicator (the same arguments on every rank); returns
    (EvalProof, v) like pcdl.open_without_eval."""
    return pcdl.open_without_eval(curve, coeffs, C, d, z, v, w, hide_q, hide_wbar, _entry=lib.halo_dist_pcdl_open)
