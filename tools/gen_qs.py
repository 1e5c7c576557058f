"""Regenerates crates/accumulation/.precompute/qs.bin in the CURRENT struct layout from GPU-produced objects -- what
`cargo run -p halo-accumulation` (crates/accumulation/src/main.rs:22-91) writes: for n = 2^2 .. 2^max a hiding
Instance::rand(n) that passes check() and the accumulator acc::prover([q]) that passes the decider, serialised as
Vec<(usize, Instance, Accumulator)>::serialize_compressed.  (The file shipped with the reference predates the current
layout and transcript and no longer deserialises; its criterion benches read this file.)

usage: tools/gen_qs.py OUT.bin [max_lg=20] [--params DIR]   DIR = crates/group/.precompute/pallas of a reference
checkout (gs-NN.bin, sh.bin); without it the reference's generator derivation is restated by the oracle."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import halo_b200 as hb


def rand_elems(rng, cnt):
    t = rng.integers(0, 1 << 63, size=(cnt, 4), dtype=np.uint64)
    t[:, 3] &= np.uint64((1 << 62) - 1)           # < 2^254 < modulus: a valid Montgomery representative of a uniform element
    return t


def generate(curve, min_lg, max_lg, seed=2025, log=print):
    rng = np.random.default_rng(seed)
    entries = []
    for lg in range(min_lg, max_lg + 1):
        n = 1 << lg
        t = time.perf_counter()
        p, z, w, wbar = rand_elems(rng, n), rand_elems(rng, 1)[0], rand_elems(rng, 1)[0], rand_elems(rng, 1)[0]
        q = hb.pcdl.Instance.open(curve, p, n - 1, z, w, rand_elems(rng, n - 1), wbar)       # Instance::rand: hiding
        q.check()
        acc = hb.acc.prover(curve, [q])
        hb.acc.decider(curve, acc)
        entries.append((n, q, acc))
        log(f"[n = 2^{lg}] instance + accumulator in {time.perf_counter() - t:.3f} s")
    return entries


if __name__ == "__main__":
    out = sys.argv[1]
    max_lg = int(sys.argv[2]) if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else 20
    hb.init()
    if "--params" in sys.argv:
        d = sys.argv[sys.argv.index("--params") + 1]
        hb.pp.new(0, 1 << max_lg, d)                 # PublicParams::new(n) from the reference's own blobs
    else:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as O                      # tool, not product: the oracle only derives the public generators
        O.set_threads(os.cpu_count() or 1)
        hb.pp.set_pp(0, *O.pp_generate(0, 1 << max_lg))
    blob = hb.serialize.qs_serialize(generate(0, 2, max_lg))
    open(out, "wb").write(blob)
    print(f"wrote {out}: {len(blob)} bytes, {max_lg - 1} entries")
