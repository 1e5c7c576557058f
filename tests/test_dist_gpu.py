"""The sharded paths behind the C ABI (halo_dist_*) on real GPUs: one process per GPU, the
library's own NCCL communicator.  Every sharded result must equal the single-GPU one AND the oracle's, bit for bit:
  MSM by point range (whole problem on every rank; ragged local slices incl. an empty one; ranks whose slices pick
  different window widths), AS_DL by instance (prover / verifier / decider, a tampered instance), NTT batches by
  polynomial (batch not divisible by the world size), one PLONK proof over all ranks (inputs on rank 0 only; both
  quotient paths; resident circuit).
World size 1 runs everywhere (pack kernel, all-gather and combine with one rank); world size 2 needs two devices."""
import os
import tempfile

import numpy as np
import pytest
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _worker(rank, world, idfile, q):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here); sys.path.insert(0, os.path.dirname(here))
    try:
        import halo_b200 as hb
        import oracle_lib as O
        hb.init(rank)
        hb.dist.init_file(rank, world, idfile)
        assert hb.dist.info() == (rank, world)
        O.set_threads(max(1, (os.cpu_count() or 2) // world))
        aff = lambda c, p: O.point_to_affine(c, np.asarray(p, dtype=np.uint64).reshape(-1, 12))
        res = {}
        for curve in (0, 1):
            sf = O.SCALAR_FIELD[curve]
            # ---- MSM: the whole problem on every rank
            k = (1 << 16) + 17
            pts = O.points_arith(curve, k, k0=0x77 + curve)
            sc = O.random_scalars(sf, k, 31 + curve)
            sc[5] = 0; pts[9] = 0
            exp = O.msm(curve, pts, sc)
            got = hb.dist.msm_global(curve, sc, pts)
            one = hb.group.point_dot_affine(curve, sc, pts)
            res[f"msm_global_{curve}"] = bool((aff(curve, got) == aff(curve, exp)).all() and (aff(curve, one) == aff(curve, exp)).all())
            # ---- ragged local slices: rank 0 owns 40000 points (window width 13), the last rank the rest (width 8 when
            # it is short), middle ranks nothing
            cut = 60000 if world > 1 else k
            lo, hi = (0, cut) if rank == 0 else (cut, k) if rank == world - 1 else (k, k)
            got = hb.dist.msm(curve, sc[lo:hi], pts[lo:hi])
            res[f"msm_ragged_{curve}"] = bool((aff(curve, got) == aff(curve, exp)).all())
            # ---- a slice large enough for the point-range-pipelined host upload (>= 2^19 points per rank)
            if curve == 0:
                kb = world * (1 << 19) + 3
                ptsb = O.points_arith(curve, kb, k0=0x99)
                scb = O.random_scalars(sf, kb, 77)
                got = hb.dist.msm_global(curve, scb, ptsb)
                res["msm_big_host_slices"] = bool((aff(curve, got) == aff(curve, O.msm(curve, ptsb, scb))).all())
            # ---- resident generators, sliced by the library
            n = 1 << 12
            gs, S, H = O.pp_generate(curve, n)
            hb.pp.reset(curve); hb.pp.set_pp(curve, gs, S, H); O.pp_set(curve, gs, S, H)
            p = O.random_scalars(sf, n, 41 + curve)
            got = hb.dist.msm_global(curve, p)
            res[f"msm_resident_{curve}"] = bool((aff(curve, got) == aff(curve, O.msm(curve, gs, p))).all())
            # ---- AS_DL by instance
            m = 11
            qs = []
            for i in range(m):
                z = O.random_scalars(sf, 1, 900 + i)[0]
                qs.append(hb.pcdl.Instance.open(curve, O.random_scalars(sf, n, 800 + i), n - 1, z))
            flat = np.stack([x.w for x in qs])
            acc1 = hb.acc.prover(curve, qs)
            accd = hb.dist.acc_prover(curve, qs)
            exp_acc = O.acc_prover(curve, flat, 12)
            fa, fb, fe = O.inst_fields(acc1.q.w), O.inst_fields(accd.q.w), O.inst_fields(exp_acc)
            same = all((fa[x] == fb[x]).all() and (fb[x] == fe[x]).all() for x in ("z", "v", "c"))
            same = same and all((aff(curve, fb[x]) == aff(curve, fe[x])).all() for x in ("C", "U", "Ls", "Rs"))
            hb.dist.acc_verifier(curve, qs, accd)
            hb.dist.acc_decider(curve, accd)
            bad = qs[m // 2].w.copy()
            bad[17:21] = O.field_add(sf, bad[17:21], O.ints_to_mont(sf, [1])[0])[0]
            qb = list(qs); qb[m // 2] = hb.pcdl.Instance(curve, bad)
            rej = 0
            for fn in (lambda: hb.dist.acc_prover(curve, qb), lambda: hb.dist.acc_verifier(curve, qb, accd),
                       lambda: hb.dist.acc_verifier(curve, qs[:-1], accd)):
                try:
                    fn()
                except hb.Rejected:
                    rej += 1
            wrongU = accd.q.w.copy()
            wrongU[21 + 24 * 12:21 + 24 * 12 + 12] = qs[0].w[0:12]                 # U := some other point of the curve
            try:
                hb.dist.acc_decider(curve, hb.acc.Accumulator(hb.pcdl.Instance(curve, wrongU)))
            except hb.Rejected:
                rej += 1
            res[f"acc_{curve}"] = bool(same and rej == 4)
            # ---- NTT batches by polynomial
            x = O.random_scalars(sf, 5 << 13, 51 + curve).reshape(5, 1 << 13, 4)
            g = O.ints_to_mont(sf, [5])[0]
            ok = (hb.dist.ntt(sf, x) == O.ntt(sf, x)).all() and (hb.dist.ntt(sf, x, inverse=True, coset=g) == O.ntt(sf, x, inverse=True, coset=g)).all()
            ok = ok and (hb.dist.ntt(sf, x[:1]) == O.ntt(sf, x[:1])).all()          # fewer polynomials than ranks
            res[f"ntt_{curve}"] = bool(ok)
            # ---- one opening over all ranks (rounds above the fold-free tail on interleaved slices): equal to the
            # single-GPU opening and, at the smaller size, to the oracle's; hiding and not; with and without the fold table
            from halo_b200._lib import lib as _l
            for lgo, check_oracle in ((17, curve == 0), (18, False)):
                no = 1 << lgo
                gso = O.points_arith(curve, no + 2, k0=0x51 + curve)
                hb.pp.reset(curve); hb.pp.set_pp(curve, gso[:no], gso[no], gso[no + 1])
                po = O.random_scalars(sf, no, 61 + curve)
                zo = O.random_scalars(sf, 4, 62 + curve)
                Co = hb.pcdl.commit(curve, po, no - 1)
                ok_open = True
                for hide in (False, True):
                    kw = dict(w=zo[1], hide_q=O.random_scalars(sf, no - 1, 63), hide_wbar=zo[2]) if hide else {}
                    Ch = hb.pcdl.commit(curve, po, no - 1, zo[1]) if hide else Co
                    for t0 in (0, 1):                                        # tuning key 0 = 1: no fold table (ladder folds)
                        _l.halo_set_tuning(0, t0)
                        pd_, vd = hb.dist.pcdl_open(curve, po, Ch, no - 1, zo[0], None, **kw)
                        p1_, v1 = hb.pcdl.open_without_eval(curve, po, Ch, no - 1, zo[0], None, **kw)
                        _l.halo_set_tuning(0, 0)
                        qd = hb.pcdl.Instance.new(curve, Ch, no - 1, zo[0], vd, pd_)
                        q1 = hb.pcdl.Instance.new(curve, Ch, no - 1, zo[0], v1, p1_)
                        fd, f1 = O.inst_fields(qd.w), O.inst_fields(q1.w)
                        ok_open = ok_open and all((fd[x] == f1[x]).all() for x in ("z", "v", "c"))
                        ok_open = ok_open and all((aff(curve, fd[x]) == aff(curve, f1[x])).all() for x in ("C", "U", "Ls", "Rs"))
                        qd.check()
                    if check_oracle and not hide:
                        O.pp_set(curve, gso[:no], gso[no], gso[no + 1])
                        fe_ = O.inst_fields(O.instance_open(curve, po, no - 1, zo[0]))
                        ok_open = ok_open and (fd["c"] == fe_["c"]).all() and (aff(curve, fd["Ls"]) == aff(curve, fe_["Ls"])).all() \
                            and (aff(curve, fd["U"]) == aff(curve, fe_["U"])).all()
                res[f"open_{curve}_{lgo}"] = bool(ok_open)
            # ---- ONE PLONK proof over all ranks: inputs on rank 0 only; equal to the oracle's proof and the single-GPU one
            # (a satisfied circuit: 8-coset quotient; a random one: 16 cosets and the coefficient-form division)
            import plonk_fixture as F
            from test_plonk_gpu import same_proof
            ok = True
            for lg, kind, npub in ((9, "gates", 0), (8, "random", 3), (3, "mul", 1)):
                rows = 1 << lg
                gsr, Sr, Hr = O.pp_generate(curve, rows)
                hb.pp.reset(curve); hb.pp.set_pp(curve, gsr, Sr, Hr); O.pp_set(curve, gsr, Sr, Hr)
                c = F.make_circuit(O, curve, lg, kind, seed=70 + lg, n_public=npub)
                acc0 = hb.acc.Accumulator.zero(curve, rows)
                circuit = hb.plonk.PlonkCircuit.from_polys(curve, c["polys"], c["n_public"])
                pub = hb.plonk.PlonkPublicInputs(c["public_inputs"], acc0)
                wit = hb.plonk.PlonkWitness(c["polys"], c["w_evals"])
                if rank == 0:
                    pd = hb.dist.plonk_prove(circuit, pub, wit)
                else:
                    pd = hb.dist.plonk_prove(circuit, int(c["public_inputs"].shape[0]), None)
                exp = O.plonk_prove(curve, lg, c["polys"], c["w_evals"], c["public_inputs"], acc0.q.w)
                same_proof(O, curve, lg, pd.w, exp)
                p1 = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
                same_proof(O, curve, lg, pd.w, p1.w)
                if kind != "random":
                    pd.verify(circuit, pub)
                # circuit columns resident on every rank, witness from rank 0
                circuit.make_resident(c["polys"])
                try:
                    pr = hb.dist.plonk_prove(circuit, pub, wit) if rank == 0 else hb.dist.plonk_prove(circuit, int(c["public_inputs"].shape[0]), None)
                finally:
                    circuit.release()
                same_proof(O, curve, lg, pr.w, exp)
            res[f"plonk_{curve}"] = ok
            hb.pp.reset(curve)
        hb.dist.shutdown()
        q.put((rank, res))
    except Exception as e:                                                       # surface the failure instead of a timeout
        import traceback
        q.put((rank, {"error": repr(e) + "\n" + traceback.format_exc()}))
        raise


def _run(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} CUDA devices")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    with tempfile.TemporaryDirectory() as d:
        idfile = os.path.join(d, "nccl_id")
        procs = [ctx.Process(target=_worker, args=(r, world, idfile, q)) for r in range(world)]
        for p in procs:
            p.start()
        out = sorted((q.get(timeout=540) for _ in procs), key=lambda t: t[0])
        for p in procs:
            p.join(timeout=60)
    for rank, res in out:
        assert "error" not in res, res["error"]
        assert res and all(res.values()), (rank, res)
    assert all(p.exitcode == 0 for p in procs)


@pytest.mark.timeout(600)
def test_dist_paths_world_size_1():
    _run(1)


@pytest.mark.timeout(600)
def test_dist_paths_world_size_2():
    _run(2)


@pytest.mark.timeout(900)
def test_dist_paths_world_size_4():
    """Four ranks: the sharded opening over four interleaved slices, and in the one-proof prover the two openings on the
    two HALVES of the communicator (ncclCommSplit)."""
    _run(4)
