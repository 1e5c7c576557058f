#!/usr/bin/env python3
"""bench.py — Pallas variable-base MSM throughput (BASELINE.json metric, config 2: random points/scalars).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--log-n 20] [--impl reference]

One "step" = one MSM over 2^log_n (default 2^20) synthetic points per GPU through libhalo_b200's C ABI.
  value  : points/s with bases and scalars already resident in HBM (halo_msm_dev), CUDA events, max over ranks
  e2e    : the same MSM through halo_msm with HOST (pinned) bases and scalars: H2D copies, kernels, D2H of the
           window sums and the host-side combine are all inside the timed region
  N > 1  : weak scaling — every rank owns a 2^log_n point range of one N*2^log_n MSM; the N partial sums
           (one Jacobian point each) are all-gathered with NCCL and added, inside the timed region
  --impl reference : the CPU restatement of the reference's path (oracle/, arkworks itself cannot be built
           here: no Rust toolchain) timed on the host cores, same workload / metric / unit.
The own arm never touches oracle/ except in the clearly separated cpu_baseline leg (rank 0, N = 1).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pallas_msm_points_per_s"
UNIT = "points/s"
PALLAS = 0
# Pallas base field (halo Fq) and scalar field (halo Fp) moduli — crates/group/src/wrappers.rs:516-517,577-578
P_BASE = 0x40000000000000000000000000000000224698fc094cf91b992d30ed00000001
Q_SCALAR = 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001
# XYZZ mixed addition = 8M + 2S (halo_b200/csrc/curve.cuh xyzz_madd); one per (point, window)
MADD_MUL, MADD_SQR = 8, 2
MODMUL_PER_MADD = MADD_MUL + MADD_SQR
# wide multiplies (IMAD.WIDE.U32) a multiplication / a squaring executes (halo_b200/csrc/field.cuh: 64 + 24, 36 + 24;
# profiles/r2_sass_histogram.txt counts them in the SASS)
IMAD_PER_MUL, IMAD_PER_SQR = 88, 60
# dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel at 2^20, from the ncu --set full
# capture of the CURRENT kernel named here; None until that capture exists (never a number from older code)
TRAFFIC_CAPTURE = {"file": "profiles/r2i_msm_accumulate_ncu.txt (ncu --set full of msm_accumulate<0>, round 2 code)",
                   "bytes_per_launch": 497.114624e6 + 64.360960e6}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--log-n", type=int, default=20)
    ap.add_argument("--impl", default="halo_b200", choices=["halo_b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-path", action="store_true", help="skip the secondary timings of the other rows of the path")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        # "under load": samples in the upper half (idle samples before/after the region sit at the idle clock)
        load = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": load[len(load) // 2] if load else None, "sm_max_mhz": mx or None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------------
# host-side helpers (python ints; verification only, outside every timed region)
# ---------------------------------------------------------------------------------------------------------
def limbs_to_int(a):
    return sum(int(x) << (64 * i) for i, x in enumerate(a))


def int_to_limbs(x):
    import numpy as np
    return np.array([(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)], dtype=np.uint64)


def jac_equal(a_xyz, b_xyz):
    """Group equality of two Jacobian points given as 12 Montgomery limbs each (Montgomery factors cancel)."""
    X1, Y1, Z1 = (limbs_to_int(a_xyz[4 * i:4 * i + 4]) for i in range(3))
    X2, Y2, Z2 = (limbs_to_int(b_xyz[4 * i:4 * i + 4]) for i in range(3))
    if Z1 == 0 or Z2 == 0:
        return Z1 == Z2
    p = P_BASE
    return (X1 * Z2 * Z2 - X2 * Z1 * Z1) % p == 0 and (Y1 * Z2 ** 3 - Y2 * Z1 ** 3) % p == 0


def gen_multipliers(seed, first, n):
    """numpy mirror of halo_gen_point_multiplier (halo_b200/csrc/capi.cu splitmix64)."""
    import numpy as np
    with np.errstate(over="ignore"):
        x = np.uint64(seed) + (np.uint64(first) + np.arange(n, dtype=np.uint64)) * np.uint64(0x2545F4914F6CDD1D)
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return x | np.uint64(1)


def expected_scalar(scalars_mont, seed, first):
    """sum_i s_i * m_i mod q for P_i = m_i * G (halo_gen_points); scalars arrive in Montgomery form.  Returns the
    canonical integer t with  MSM == t * G."""
    s = scalars_mont.astype(object)
    vals = s[:, 0] + (s[:, 1] << 64) + (s[:, 2] << 128) + (s[:, 3] << 192)
    m = gen_multipliers(seed, first, scalars_mont.shape[0]).astype(object)
    acc = int((vals * m).sum())
    return (acc % Q_SCALAR) * pow(1 << 256, -1, Q_SCALAR) % Q_SCALAR


def cpu_msm_leg(log_n, reps):
    """Times the CPU restatement (oracle/msm.hpp, all host threads) on a 2^log_n Pallas MSM.  The ONLY place
    bench.py executes anything under oracle/."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cores = os.cpu_count() or 1
    O.set_threads(cores)
    k = 1 << log_n
    pts = O.points_arith(PALLAS, k)
    sc = O.random_scalars(O.SCALAR_FIELD[PALLAS], k, 20)
    times = []
    for _ in range(reps):
        t = time.perf_counter()
        O.msm(PALLAS, pts, sc)
        times.append(time.perf_counter() - t)
    return k, cores, times


def run_reference(args, rank):
    if rank != 0:
        return
    # bounded sample: the full 2^20 MSM is ~10-20 s of single-core work, a few seconds on all cores
    log_n = min(args.log_n, 20)
    k, cores, times = cpu_msm_leg(log_n, args.warmup + args.steps)
    timed = times[args.warmup:]
    total = sum(timed)
    v = k * len(timed) / total
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(timed), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32x8-montgomery", "data": "synthetic",
        "config": {"workload": f"pallas_msm_2^{log_n}_random_points", "log_n": log_n,
                   "note": "CPU restatement of the reference path (oracle/), not arkworks: no Rust toolchain on this box"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{len(timed)} x MSM of 2^{log_n} points, all host threads"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def measure_path(lib, check, halo_b200, torch, stream, d_pts, h_sc, log_n):
    """Secondary timings of the other rows of the hot path at the same size (reported, not the headline): NTT / iNTT,
    batched commit, IPA opening, succinct check and full check.  Device-resident where the ABI allows it."""
    import numpy as np
    n = 1 << log_n
    u64p = ctypes.POINTER(ctypes.c_uint64)
    out = {}

    def ev_ms(fn, reps=3):
        best = None
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); fn(); e1.record(stream); e1.synchronize()
            t = e0.elapsed_time(e1)
            best = t if best is None else min(best, t)
        return best

    for b in (1, 16):
        x = torch.randint(-(1 << 63), (1 << 63) - 1, (b * n, 4), dtype=torch.int64, device="cuda")
        x[:, 3] &= (1 << 62) - 1
        f = lambda: check(lib.halo_ntt_dev(0, ctypes.c_void_p(x.data_ptr()), ctypes.c_uint(log_n), ctypes.c_size_t(b), 0, None))
        f()
        ms = ev_ms(f)
        passes = 1 + max(0, -(-(log_n - 11) // 9))           # ntt.cu: a last pass of <= 11 stages + passes of <= 9
        out[f"ntt_2^{log_n}_x{b}"] = {"ms": ms, "elements_per_s": b * n / (ms * 1e-3),
                                      "gbutterfly_per_s": b * n * log_n / 2 / (ms * 1e-3) / 1e9,
                                      "passes": passes,
                                      "hbm_gb_per_s": 64.0 * b * n * passes / (ms * 1e-3) / 1e9}   # 64 B per element and pass
        del x
    # the points of the MSM benchmark double as generators; S and H are two more of them
    gs = d_pts.cpu().numpy().view(np.uint64)
    sh = np.zeros((2, 8), dtype=np.uint64)
    check(lib.halo_gen_points(PALLAS, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
    halo_b200.pp.reset(PALLAS)
    halo_b200.pp.set_pp(PALLAS, gs, sh[0], sh[1])
    p = h_sc.numpy().view(np.uint64)
    z = p[7].copy()
    if n <= (1 << 20):
        # the 16 commits of a prover round (protocol.rs:114,263) as one pipeline over the resident generators
        b = 16
        cols = torch.randint(-(1 << 63), (1 << 63) - 1, (b * n, 4), dtype=torch.int64, device="cuda")
        cols[:, 3] &= (1 << 62) - 1
        res = np.zeros((b, 12), dtype=np.uint64)
        f = lambda: check(lib.halo_msm_batch_dev(PALLAS, None, ctypes.c_size_t(0), ctypes.c_void_p(cols.data_ptr()),
                                                 ctypes.c_size_t(n), ctypes.c_size_t(b), res.ctypes.data_as(u64p)))
        f()
        ms = ev_ms(f, reps=2)
        out[f"commit_batch16_2^{log_n}"] = {"ms": ms, "points_per_s": b * n / (ms * 1e-3)}
        del cols
    C = halo_b200.pcdl.commit(PALLAS, p, n - 1)
    best = None
    for _ in range(2):
        t = time.perf_counter(); pi, v = halo_b200.pcdl.open_without_eval(PALLAS, p, C, n - 1, z, None); dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    out[f"pcdl_open_2^{log_n}_e2e"] = {"ms": best * 1e3}
    q = halo_b200.pcdl.Instance.new(PALLAS, C, n - 1, z, v, pi)
    t = time.perf_counter(); q.succinct_check(); out["pcdl_succinct_check"] = {"ms": (time.perf_counter() - t) * 1e3}
    t = time.perf_counter(); q.check(); out[f"pcdl_check_2^{log_n}"] = {"ms": (time.perf_counter() - t) * 1e3}
    if log_n > 14:                                   # the size the CPU leg opens (bounded sample)
        ps = p[: 1 << 14]
        Cs = halo_b200.pcdl.commit(PALLAS, ps, (1 << 14) - 1)
        t = time.perf_counter(); halo_b200.pcdl.open_without_eval(PALLAS, ps, Cs, (1 << 14) - 1, z, None)
        out["pcdl_open_2^14_e2e"] = {"ms": (time.perf_counter() - t) * 1e3}
    # BASELINE.json's second headline: PLONK proof time at 2^log_n gates (synthetic
    # multiplication circuit, genesis accumulator), through the public API with host buffers, plus the same size
    # with every custom gate's selector non-zero, and 2^12 for the like-for-like CPU ratio
    for lg in sorted({12, min(log_n, 20)}):
        rows = 1 << lg
        acc0 = halo_b200.acc.Accumulator.zero(PALLAS, rows)
        pub = halo_b200.plonk.PlonkPublicInputs(np.zeros((0, 4), dtype=np.uint64), acc0)
        # three circuits: the config-4 multiplication circuit; every selector / witness / round-constant column random with
        # the constant selector solved so that the circuit IS satisfied (what an honest prover proves: 8-coset quotient);
        # the same with an unsatisfied witness (the 16-coset fallback the reference's divide_by_vanishing_poly implies)
        for dense, sat in ((False, False), (True, True), (True, False)):
            polys, w_evals = halo_b200.plonk.synthetic_circuit(PALLAS, lg, seed=7, dense_gates=dense, satisfied=sat)
            circuit = halo_b200.plonk.PlonkCircuit.from_polys(PALLAS, polys, 0)
            wit = halo_b200.plonk.PlonkWitness(polys, w_evals)
            best = None
            for _ in range(2):
                t = time.perf_counter(); pi = halo_b200.plonk.PlonkProof.naive_prover(circuit, pub, wit); dt = time.perf_counter() - t
                best = dt if best is None else min(best, dt)
            key = f"plonk_prove_2^{lg}" + (("_all_gates" if sat else "_all_gates_unsatisfied_witness") if dense else "")
            out[key] = {"ms": best * 1e3, "h2d_bytes": int(polys.nbytes + w_evals.nbytes)}
            if dense and sat:
                t = time.perf_counter(); pi.verify(circuit, pub); out[key]["verify_ms"] = (time.perf_counter() - t) * 1e3
            if not dense:
                t = time.perf_counter(); pi.verify(circuit, pub); out[f"plonk_verify_2^{lg}"] = {"ms": (time.perf_counter() - t) * 1e3}
                # the circuit's own 41 columns resident in HBM (halo_plonk_circuit_load): a proof uploads its witness only
                circuit.make_resident(polys)
                best = None
                for _ in range(2):
                    t = time.perf_counter(); pi2 = halo_b200.plonk.PlonkProof.naive_prover(circuit, pub, wit); dt = time.perf_counter() - t
                    best = dt if best is None else min(best, dt)
                pi2.verify(circuit, pub)
                out[key + "_circuit_resident"] = {"ms": best * 1e3, "h2d_bytes": int(polys[10:26].nbytes + w_evals.nbytes)}
                circuit.release()
            del polys, w_evals, wit
    halo_b200.pp.reset(PALLAS)
    return out


def cpu_path_leg():
    """CPU restatement timings for the other rows (bounded samples): NTT 2^20 and a PC_DL opening at 2^14."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_lib as O
    O.set_threads(os.cpu_count() or 1)
    out = {}
    x = O.random_scalars(0, 1 << 20, 3)
    t = time.perf_counter(); O.ntt(0, x); out["ntt_2^20_ms"] = (time.perf_counter() - t) * 1e3
    n = 1 << 14
    gs = O.points_arith(PALLAS, n + 2)
    O.pp_set(PALLAS, gs[:n], gs[n], gs[n + 1])
    p = O.random_scalars(0, n, 4)
    t = time.perf_counter(); O.instance_open(PALLAS, p, n - 1, p[3]); out["pcdl_instance_open_2^14_ms"] = (time.perf_counter() - t) * 1e3
    # the PLONK prover restatement on a 2^12-row multiplication circuit (bounded sample of config 4)
    import plonk_fixture as F
    lg = 12
    rows = 1 << lg
    c = F.make_circuit(O, PALLAS, lg, "mul", seed=7)
    zero = O.instance_open(PALLAS, np.zeros((0, 4), dtype=np.uint64), rows - 1, np.zeros(4, dtype=np.uint64))
    acc0 = O.acc_prover(PALLAS, zero.reshape(1, -1), lg)
    t = time.perf_counter(); O.plonk_prove(PALLAS, lg, c["polys"], c["w_evals"], c["public_inputs"], acc0)
    out["plonk_prove_2^12_ms"] = (time.perf_counter() - t) * 1e3
    return out


def run_halo(args, rank, world, local_rank):
    import numpy as np
    import torch
    import halo_b200
    from halo_b200._lib import lib, check

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — halo_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    halo_b200.init(local_rank)
    if world > 1:
        halo_b200.dist.init_torch(dist)      # the library's own NCCL communicator (halo_dist_init); id over torch.distributed
    stream = torch.cuda.Stream()      # torch's default stream has handle 0, which halo_set_stream reads as 'library stream'
    torch.cuda.set_stream(stream)
    check(lib.halo_set_stream(ctypes.c_void_p(stream.cuda_stream)))
    lib.halo_launch_count.restype = ctypes.c_ulonglong
    u64p = ctypes.POINTER(ctypes.c_uint64)

    k = 1 << args.log_n
    seed, first = 0x48414C4F, rank * k
    # ---- synthetic inputs: distinct points P_i = m_i*G generated on the device, uniform scalars (any 254-bit
    # value is a valid Montgomery representative of a uniformly distributed field element)
    d_pts = torch.empty((k, 8), dtype=torch.int64, device="cuda")
    check(lib.halo_gen_points_dev(PALLAS, ctypes.c_uint64(seed), ctypes.c_size_t(first), ctypes.c_size_t(k),
                                  ctypes.c_void_p(d_pts.data_ptr())))
    g = torch.Generator(device="cuda"); g.manual_seed(1234 + rank)
    d_sc = torch.randint(-(1 << 63), (1 << 63) - 1, (k, 4), dtype=torch.int64, device="cuda", generator=g)
    d_sc[:, 3] &= (1 << 62) - 1
    h_pts = d_pts.cpu().pin_memory()
    h_sc = d_sc.cpu().pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # 256 MiB > 126 MB L2
    total = np.zeros(12, dtype=np.uint64)

    prof = {"on": False, "ms": {}}
    ms_buf = (ctypes.c_float * 24)()
    nm_buf = (ctypes.c_char_p * 24)()

    def read_stage_profile():
        if prof["on"]:                       # per-stage CUDA-event times of THIS MSM
            nst = lib.halo_profile_read(ms_buf, nm_buf, 24)
            for j in range(max(nst, 0)):
                nm = nm_buf[j].decode()
                prof["ms"][nm] = prof["ms"].get(nm, 0.0) + ms_buf[j]

    # N > 1: one call per rank does the whole point-range-sharded MSM inside the library: local Pippenger down to the
    # window sums, ncclAllGather of them straight from device memory, combine on every rank (halo_dist_msm[_dev])
    def step_dev():
        if world == 1:
            check(lib.halo_msm_dev(PALLAS, ctypes.c_void_p(d_pts.data_ptr()), ctypes.c_size_t(0),
                                   ctypes.c_void_p(d_sc.data_ptr()), ctypes.c_size_t(k), total.ctypes.data_as(u64p)))
        else:
            check(lib.halo_dist_msm_dev(PALLAS, ctypes.c_void_p(d_pts.data_ptr()), ctypes.c_size_t(0),
                                        ctypes.c_void_p(d_sc.data_ptr()), ctypes.c_size_t(k), total.ctypes.data_as(u64p)))
        read_stage_profile()

    def make_step_host(pts_arr_ptr, sc_arr_ptr):
        def f():
            fn = lib.halo_msm if world == 1 else lib.halo_dist_msm
            check(fn(PALLAS, pts_arr_ptr, ctypes.c_size_t(0), sc_arr_ptr, ctypes.c_size_t(k), total.ctypes.data_as(u64p)))
        return f

    step_e2e = make_step_host(ctypes.cast(h_pts.data_ptr(), u64p), ctypes.cast(h_sc.data_ptr(), u64p))
    # the same through PAGEABLE host buffers (what a Rust Vec is): the library stages them through pinned chunks
    pg_pts, pg_sc = h_pts.numpy().copy(), h_sc.numpy().copy()
    step_e2e_pageable = make_step_host(ctypes.cast(pg_pts.ctypes.data, u64p), ctypes.cast(pg_sc.ctypes.data, u64p))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, collect_profile=False):
        """K steps, each bracketed by CUDA events on the launching stream with an L2 flush in between (outside the
        events); returns total ms (max over ranks) and per-stage sums."""
        prof["on"], prof["ms"] = collect_profile, {}
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        for i in range(steps):
            flush.zero_()
            ev[i][0].record(stream)
            fn()
            ev[i][1].record(stream)
        barrier()
        prof["on"] = False
        stage_ms = dict(prof["ms"])
        tot = sum(a.elapsed_time(b) for a, b in ev)
        if dist is not None:
            t = torch.tensor([tot], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            tot = float(t.item())
        return tot, stage_ms

    # ---- measured integer-pipe peak: dependent Montgomery multiplications, 4 chains/thread (capi.cu) ----
    ms = ctypes.c_float()
    chk = (ctypes.c_uint64 * 4)()
    peak_modmul = 0.0
    for threads, bps in ((256, 4), (512, 2)):
        check(lib.halo_bench_modmul(1, 148 * bps, threads, 400, ctypes.byref(ms), chk))
        peak_modmul = max(peak_modmul, 148 * bps * threads * 400 * 4 / (ms.value * 1e-3))

    peak_imad = 0.0
    chk1 = (ctypes.c_uint64 * 1)()
    for threads, bps in ((256, 4), (512, 2)):
        check(lib.halo_bench_imad(148 * bps, threads, 4000, ctypes.byref(ms), chk1))
        peak_imad = max(peak_imad, 148 * bps * threads * 4000 * 8 / (ms.value * 1e-3))

    # ---- device-resident arm ----
    for _ in range(max(args.warmup, 3)):
        step_dev()
    check(lib.halo_profile_enable(1))
    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = lib.halo_launch_count()
    tot_ms, stage_ms = timed(step_dev, args.steps, collect_profile=True)
    launches = lib.halo_launch_count() - l0
    check(lib.halo_profile_enable(0))
    result_dev = total.copy()

    # ---- end-to-end arm (host buffers) ----
    for _ in range(max(args.warmup, 3)):
        step_e2e()
    e2e_ms, _ = timed(step_e2e, args.steps)
    result_e2e = total.copy()
    for _ in range(3):
        step_e2e_pageable()
    e2e_pg_ms, _ = timed(step_e2e_pageable, args.steps)
    clocks = sampler.stop() if sampler else None
    result_pg = total.copy()

    # ---- self-consistency check against the known discrete logs (outside the timed regions).  The points are m_i * G, so
    # the MSM must equal (sum s_i m_i) * G; the right-hand side is a k = 1 MSM of THIS library, hence "self_check": the
    # independent bit-exact check of a 2^20-point MSM against the oracle is tests/test_msm_gpu.py::test_msm_large_bit_exact
    # (and the cpu_baseline leg below compares this very result with the oracle's scalar multiplication at N = 1).
    def all_sum(t_loc):
        if dist is None:
            return t_loc
        parts = [None] * world
        dist.all_gather_object(parts, t_loc)
        return sum(parts) % Q_SCALAR

    def point_of_scalar(t):
        """t * G through a k = 1 MSM of the library (gen[0] = m0 * G  =>  t*G = (t / m0) * gen[0])."""
        exp = np.zeros(12, dtype=np.uint64)
        gen = np.zeros((1, 8), dtype=np.uint64)
        check(lib.halo_gen_points(PALLAS, ctypes.c_uint64(0), ctypes.c_size_t(0), ctypes.c_size_t(1), gen.ctypes.data_as(u64p)))
        m0 = int(gen_multipliers(0, 0, 1)[0])
        t_adj = t * pow(m0, -1, Q_SCALAR) % Q_SCALAR
        sc1 = int_to_limbs(t_adj * (1 << 256) % Q_SCALAR).reshape(1, 4)
        check(lib.halo_msm(PALLAS, gen.ctypes.data_as(u64p), ctypes.c_size_t(0), sc1.ctypes.data_as(u64p),
                           ctypes.c_size_t(1), exp.ctypes.data_as(u64p)))
        return exp

    t_all = all_sum(expected_scalar(d_sc.cpu().numpy().view(np.uint64), seed, first))
    self_check = None
    if rank == 0:
        exp = point_of_scalar(t_all)
        self_check = bool(jac_equal(result_dev, exp) and jac_equal(result_e2e, exp) and jac_equal(result_pg, exp))
        if not self_check:
            raise SystemExit("bench.py: MSM result does not match the known discrete-log check")

    # ---- STRONG scaling of the headline: ONE MSM of a fixed total size (2^20 and 2^22 points) split by point range over
    # the N GPUs, device-resident, exchange inside the timed region, CUDA events, max over ranks ----
    strong = {}
    if not args.no_path:
        for lgT in (20, 22):
            kT = 1 << lgT
            lo, hi = kT * rank // world, kT * (rank + 1) // world
            kl = hi - lo
            s_pts = torch.empty((kl, 8), dtype=torch.int64, device="cuda")
            check(lib.halo_gen_points_dev(PALLAS, ctypes.c_uint64(seed + 1), ctypes.c_size_t(lo), ctypes.c_size_t(kl),
                                          ctypes.c_void_p(s_pts.data_ptr())))
            gs_ = torch.Generator(device="cuda"); gs_.manual_seed(999 + 31 * lgT + rank)
            s_sc = torch.randint(-(1 << 63), (1 << 63) - 1, (kl, 4), dtype=torch.int64, device="cuda", generator=gs_)
            s_sc[:, 3] &= (1 << 62) - 1
            res = np.zeros(12, dtype=np.uint64)

            def step_strong():
                fn = lib.halo_msm_dev if world == 1 else lib.halo_dist_msm_dev
                check(fn(PALLAS, ctypes.c_void_p(s_pts.data_ptr()), ctypes.c_size_t(0), ctypes.c_void_p(s_sc.data_ptr()),
                         ctypes.c_size_t(kl), res.ctypes.data_as(u64p)))
            for _ in range(3):
                step_strong()
            reps = max(5, args.steps // 2)
            ms_tot, _ = timed(step_strong, reps)
            t_s = all_sum(expected_scalar(s_sc.cpu().numpy().view(np.uint64), seed + 1, lo))
            ok = bool(jac_equal(res, point_of_scalar(t_s))) if rank == 0 else None
            if rank == 0 and not ok:
                raise SystemExit(f"bench.py: strong-scaling MSM 2^{lgT} fails the discrete-log check")
            strong[f"msm_2^{lgT}_total"] = {"ms_per_step": ms_tot / reps, "points_per_s": kT * reps / (ms_tot * 1e-3),
                                            "points_per_gpu": kl, "self_check": ok}
            del s_pts, s_sc

    # ---- config 5 (BASELINE.json): AS_DL prove + verify + decide over m = 1024 instances of n = 2^16, sharded by instance
    # at N > 1 (halo_dist_acc_*); wall clock, max over ranks.  The transcript absorbs every challenge of every
    # instance in ONE serial sponge (acc.rs:157-160), which bounds what sharding can give.
    acc5 = None
    if not args.no_path:
        lg5, m5 = 16, 1024
        n5 = 1 << lg5
        g5 = np.zeros((n5 + 2, 8), dtype=np.uint64)
        check(lib.halo_gen_points(PALLAS, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(n5 + 2), g5.ctypes.data_as(u64p)))
        halo_b200.pp.reset(PALLAS)
        halo_b200.pp.set_pp(PALLAS, g5[:n5], g5[n5], g5[n5 + 1])
        rng5 = np.random.default_rng(5)                    # same seed on every rank: identical instances

        def rand_elems(cnt):
            t = rng5.integers(0, 1 << 63, size=(cnt, 4), dtype=np.uint64)
            t[:, 3] &= np.uint64((1 << 62) - 1)
            return t
        base5 = [halo_b200.pcdl.Instance.open(PALLAS, rand_elems(n5), n5 - 1, rand_elems(1)[0]) for _ in range(32)]
        qs5 = [base5[i % 32] for i in range(m5)]
        P5 = (halo_b200.acc.prover, halo_b200.acc.verifier, halo_b200.acc.decider) if world == 1 else \
             (halo_b200.dist.acc_prover, halo_b200.dist.acc_verifier, halo_b200.dist.acc_decider)
        best = {}

        def wall(name, fn):
            barrier()
            t0 = time.perf_counter(); r = fn(); dt = time.perf_counter() - t0
            if dist is not None:
                tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                dt = float(tt.item())
            best[name] = min(best.get(name, 1e9), dt)
            return r
        for _ in range(2):
            acc_ = wall("prover", lambda: P5[0](PALLAS, qs5))
            wall("verifier", lambda: P5[1](PALLAS, qs5, acc_))
            wall("decider", lambda: P5[2](PALLAS, acc_))
        acc5 = {"m": m5, "log_n": lg5, "prover_ms": best["prover"] * 1e3, "verifier_ms": best["verifier"] * 1e3,
                "decider_ms": best["decider"] * 1e3, "sharding": "by instance (halo_dist_acc_*)" if world > 1 else "single GPU",
                "note": "dominated by the protocol's serial Poseidon sponge over all m*(lg n + 1) challenges (host)"}
        halo_b200.pp.reset(PALLAS)

    # whole proofs at N > 1, two ways: replicas (every rank proves its own 2^log_n-gate circuit at the same time; the
    # time is the max over ranks) and ONE proof sharded over all ranks (further down)
    replicas = None
    if world > 1 and not args.no_path:
        lg = min(args.log_n, 20)
        rows = 1 << lg
        sh = np.zeros((2, 8), dtype=np.uint64)
        check(lib.halo_gen_points(PALLAS, ctypes.c_uint64(5), ctypes.c_size_t(0), ctypes.c_size_t(2), sh.ctypes.data_as(u64p)))
        halo_b200.pp.reset(PALLAS)
        halo_b200.pp.set_pp(PALLAS, h_pts.numpy().view(np.uint64)[:rows], sh[0], sh[1])
        polys, w_evals = halo_b200.plonk.synthetic_circuit(PALLAS, lg, seed=7 + rank)
        acc0 = halo_b200.acc.Accumulator.zero(PALLAS, rows)
        circuit = halo_b200.plonk.PlonkCircuit.from_polys(PALLAS, polys, 0)
        pub = halo_b200.plonk.PlonkPublicInputs(np.zeros((0, 4), dtype=np.uint64), acc0)
        wit = halo_b200.plonk.PlonkWitness(polys, w_evals)
        def timed_proofs():
            best = None
            for _ in range(2):
                dist.barrier(); torch.cuda.synchronize()
                t = time.perf_counter(); pi = halo_b200.plonk.PlonkProof.naive_prover(circuit, pub, wit); dt = time.perf_counter() - t
                tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                best = float(tt.item()) if best is None else min(best, float(tt.item()))
            pi.verify(circuit, pub)
            return best
        best = timed_proofs()
        circuit.make_resident(polys)                        # the circuit's own columns stay in HBM; witness-only upload
        best_res = timed_proofs()
        circuit.release()
        replicas = {"ms": best * 1e3, "proofs_per_s": world / best, "ms_circuit_resident": best_res * 1e3,
                    "proofs_per_s_circuit_resident": world / best_res,
                    "note": "one independent proof per GPU, max over ranks"}
        # ONE proof over all N GPUs (halo_dist_plonk_prove: inputs on rank 0, columns broadcast over NVLink, commitments as
        # point-range-sharded MSMs, quotient cosets divided among the ranks, the two openings on ranks 0 and 1); the same
        # circuit on every rank so that the resident variant can be timed too.  Parity unpinned like every PLONK figure.
        one = {}
        # the sharded proof needs the SAME generators on every rank (h_pts is this rank's slice of the weak-scaling MSM)
        gp = torch.empty((rows, 8), dtype=torch.int64, device="cuda")
        check(lib.halo_gen_points_dev(PALLAS, ctypes.c_uint64(99), ctypes.c_size_t(0), ctypes.c_size_t(rows), ctypes.c_void_p(gp.data_ptr())))
        halo_b200.pp.reset(PALLAS)
        halo_b200.pp.set_pp(PALLAS, gp.cpu().numpy().view(np.uint64), sh[0], sh[1])
        del gp
        acc0 = halo_b200.acc.Accumulator.zero(PALLAS, rows)            # the genesis accumulator belongs to the generator set
        pub = halo_b200.plonk.PlonkPublicInputs(np.zeros((0, 4), dtype=np.uint64), acc0)
        for dense in (False, True):
            polys, w_evals = halo_b200.plonk.synthetic_circuit(PALLAS, lg, seed=7, dense_gates=dense, satisfied=dense)
            circuit = halo_b200.plonk.PlonkCircuit.from_polys(PALLAS, polys, 0)
            wit = halo_b200.plonk.PlonkWitness(polys, w_evals)
            def timed_one():
                best = None
                for _ in range(3):
                    dist.barrier(); torch.cuda.synchronize()
                    t = time.perf_counter()
                    pi = halo_b200.dist.plonk_prove(circuit, pub, wit) if rank == 0 else halo_b200.dist.plonk_prove(circuit, 0, None)
                    dt = time.perf_counter() - t
                    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
                    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                    best = float(tt.item()) if best is None else min(best, float(tt.item()))
                return best, pi
            t_one, pi = timed_one()
            pi.verify(circuit, pub)                       # both circuits are satisfied: the proofs must be accepted
            key = "all_gates" if dense else "mul_circuit"
            one[key + "_ms"] = t_one * 1e3
            if not dense:
                circuit.make_resident(polys)
                t_res, _ = timed_one()
                circuit.release()
                one[key + "_circuit_resident_ms"] = t_res * 1e3
            del polys, w_evals, wit
        one["note"] = ("ONE 2^%d-gate proof sharded over %d GPUs behind the C ABI (halo_dist_plonk_prove), host buffers on rank 0, "
                       "max over ranks; parity unpinned (no reference proof exists)" % (lg, world))
        replicas["one_proof"] = one
        halo_b200.pp.reset(PALLAS)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    n_pts = k * world
    value = n_pts * args.steps / (tot_ms * 1e-3)
    e2e_value = n_pts * args.steps / (e2e_ms * 1e-3)
    acc_ms = stage_ms.get("accumulate", 0.0) / args.steps
    lg = args.log_n
    c = 16 if lg >= 20 else 15 if lg >= 17 else 13 if lg >= 15 else 8; nw = 255 // c + 1     # msm.cu pick_window / make_plan
    roof = None
    if acc_ms > 0:
        # dominant kernel: bucket accumulation.  Algorithmic work per launch: one mixed addition (10 modmul) per
        # (point, window); algorithmic bytes per launch: 4 B index + 64 B base per (point, window) + bucket writes
        modmul = MODMUL_PER_MADD * k * nw
        abytes = k * nw * (4 + 64) + (nw << (c - 1)) * 128
        ach = modmul / (acc_ms * 1e-3)
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        step_s = tot_ms / args.steps * 1e-3
        # SURVEY section 8(d)'s canonical model of the WHOLE step (signed 16-bit windows, 11-modmul mixed additions, full
        # bucket reduction): W_msm(n) = 176 n + 1.68e7 modmul; work per rank
        model_modmul = 176.0 * k + 1.68e7
        # executed wide multiplies of the dominant kernel: squarings are 60, not 88
        imads = (MADD_MUL * IMAD_PER_MUL + MADD_SQR * IMAD_PER_SQR) * k * nw
        roof = {"kernel": "msm_accumulate", "bound": "int32-pipe", "achieved": ach / 1e9, "peak": peak_modmul / 1e9,
                "unit": "Gmodmul/s", "frac": ach / peak_modmul,
                "traffic": TRAFFIC_CAPTURE["bytes_per_launch"] if args.log_n == 20 else None,
                "traffic_source": TRAFFIC_CAPTURE["file"],
                "peak_source": "measured live: halo_bench_modmul (dependent 255-bit Montgomery multiplications of THIS library, "
                               "i.e. a ceiling for code built on fe_mul, not the raw pipe: see `pipe`)",
                "ms_per_launch": acc_ms, "share_of_step": acc_ms / (tot_ms / args.steps),
                "pipe": {"what": "executed IMAD.WIDE.U32 of the kernel (8 x 88 + 2 x 60 per mixed addition) / raw issue rate",
                         "achieved": imads / (acc_ms * 1e-3) / 1e12, "peak": peak_imad / 1e12, "unit": "T IMAD.WIDE/s",
                         "frac": imads / (acc_ms * 1e-3) / peak_imad,
                         "peak_source": "measured live: halo_bench_imad (8 independent wide multiply-adds per thread)"},
                "step_frac": {"what": "SURVEY 8(d) model 176 n + 1.68e7 modmul over the whole step / live modmul peak",
                              "achieved": model_modmul / step_s / 1e9, "peak": peak_modmul / 1e9, "unit": "Gmodmul/s",
                              "frac": model_modmul / step_s / peak_modmul},
                "hbm": {"bound": "hbm", "achieved": abytes / (acc_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": abytes / (acc_ms * 1e-3) / 1e9 / hbm_peak,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
                "stages_ms": {kk: v / args.steps for kk, v in stage_ms.items()}}

    path = None
    if world == 1 and not args.no_path:
        path = measure_path(lib, check, halo_b200, torch, stream, d_pts, h_sc, args.log_n)
    elif replicas is not None:
        path = {f"plonk_prove_2^{min(args.log_n, 20)}_replicas": replicas}
    if path is not None:
        path["strong_scaling"] = strong
        path[f"asdl_m1024_2^16"] = acc5
        path["parity_note"] = ("PLONK figures: parity unpinned (the reference stores no proof; GPU == CPU restatement field by "
                               "field up to 2^16 rows, tests/test_config_sizes_gpu.py); MSM / IPA / AS_DL / NTT: pinned")

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        kk, cores, times = cpu_msm_leg(min(args.log_n, 20), 2)
        # the checker, in its own leg: t * G by the oracle's scalar multiplication must be the point the timed GPU runs returned
        import oracle_lib as O_
        gen_aff = O_.generator(PALLAS)
        exp_o = O_.point_mul(PALLAS, gen_aff, int_to_limbs(t_all * (1 << 256) % Q_SCALAR))
        oracle_check = bool(jac_equal(result_dev, np.asarray(exp_o, dtype=np.uint64).reshape(-1)[:12]))
        if not oracle_check:
            raise SystemExit("bench.py: the timed MSM result differs from the oracle's t * G")
        cpu = {"value": kk / min(times), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"best of 2 x MSM of 2^{min(args.log_n, 20)} points, CPU restatement (oracle/msm.hpp), all host threads"}
        cpu["timed_result_equals_oracle"] = oracle_check
        if not args.no_path:
            cpu["path"] = cpu_path_leg()

    print(json.dumps({
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": tot_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32x8-montgomery", "data": "synthetic",
        "config": {"workload": f"pallas_msm_2^{args.log_n}_random_points", "log_n": args.log_n, "points_per_gpu": k,
                   "sharding": "point range; window sums all-gathered by the library (halo_dist_msm_dev)" if world > 1 else "single GPU",
                   "l2": "flushed between timed iterations (256 MiB memset outside the event pair)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(k * 96), "d2h_bytes_per_step": int(nw * 128),
                "ms_per_step": e2e_ms / args.steps, "host_buffers": "pinned",
                "pageable": {"value": n_pts * args.steps / (e2e_pg_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_pg_ms / args.steps,
                             "note": "plain malloc'd buffers (a Rust Vec): staged through pinned chunks (non-temporal copies) by the library's host threads"}},
        "gpu_launches": int(launches), "self_check": self_check, "clocks": clocks, "roofline": roof, "cpu_baseline": cpu, "path": path,
    }))
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-launch one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    run_halo(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
