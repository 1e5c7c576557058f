"""Synthetic circuits in the shape Trace::new hands to the prover (crates/plonk/src/circuit/trace.rs:137-232): column
vectors -> Evals (row i at omega^(i+1)) -> coefficient vectors, plus the identity/sigma columns of build_sigma
(trace.rs:64-104).  Pure Python integers and the oracle's iNTT; test infrastructure."""
import random

import numpy as np

Q, W, R_, S = 10, 16, 15, 8


def _poseidon_mds(O, field):
    import re, os
    txt = open(os.path.join(os.path.dirname(os.path.abspath(O.__file__)), "..", "oracle", "poseidon_consts.inc")).read()
    body = txt[txt.index("POSEIDON_MDS"):txt.index("POSEIDON_RC")]
    vals = [int(x, 16) for x in re.findall(r"0x([0-9a-fA-F]+)ULL", body)]
    limbs = np.array(vals, dtype=np.uint64).reshape(2, 9, 4)
    return [O.from_mont(field, limbs[field][i]) for i in range(9)]


def columns_to_polys(O, field, cols, lg_n):
    """cols: list of n-int columns -> (k, n, 4) coefficient arrays and (k, n, 4) Evals-convention value arrays."""
    n = 1 << lg_n
    ev = np.stack([np.roll(O.ints_to_mont(field, c), 1, axis=0) for c in cols])          # from_vec_and_domain
    co = O.ntt(field, ev.copy(), inverse=True)
    return co.reshape(len(cols), n, 4), ev


def make_circuit(O, curve, lg_n, kind="mul", seed=1, n_public=0):
    """Returns dict(polys (57,n,4), w_evals (3,n,4), public_inputs (k,4), cols) for the scalar field of `curve`."""
    field = curve            # a curve's scalar field id equals its curve id
    p = O.MODULUS[field]
    n = 1 << lg_n
    rng = random.Random(seed * 7919 + lg_n)
    qs = [[0] * n for _ in range(Q)]
    ws = [[0] * n for _ in range(W)]
    rs = [[0] * n for _ in range(R_)]
    ids = [[(j + 1) + i * n for j in range(n)] for i in range(S)]
    sig = [list(c) for c in ids]
    pub = []
    rnd = lambda: rng.randrange(p)
    if kind == "random":                     # not satisfiable: exercises every gate expression of the prover
        qs = [[rnd() for _ in range(n)] for _ in range(Q)]
        ws = [[rnd() for _ in range(n)] for _ in range(W)]
        rs = [[rnd() for _ in range(n)] for _ in range(R_)]
        sig = [[rnd() for _ in range(n)] for _ in range(S)]
        pub = [rnd() for _ in range(n_public)]
    elif kind in ("mul", "copy"):            # a*b - c = 0 on every row
        for j in range(n):
            a, b = rnd(), rnd()
            ws[0][j], ws[1][j], ws[2][j] = a, b, a * b % p
            qs[3][j], qs[2][j] = 1, p - 1
        if kind == "copy":                   # tie row j+1's a to row j's c for even j: 2-cycles in sigma
            for j in range(0, n - 1, 2):
                ws[0][j + 1] = ws[2][j]
                ws[2][j + 1] = ws[0][j + 1] * ws[1][j + 1] % p
                sig[0][j + 1], sig[2][j] = ids[2][j], ids[0][j + 1]
        assert kind == "mul" or n_public == 0
        for k in range(n_public):            # public input rows: q_l * a + PI = 0 with PI = -x
            x = rnd()
            pub.append(x)
            ws[0][k], ws[1][k], ws[2][k] = x, 0, 0
            for q in qs:
                q[k] = 0
            qs[0][k] = 1
    elif kind == "gates":                    # satisfiable rows for the custom gates that need no curve points
        M = _poseidon_mds(O, field)
        j = 0
        while j + 2 <= n:
            g = rng.randrange(4)
            if g == 0:                       # multiplication
                a, b = rnd(), rnd()
                ws[0][j], ws[1][j], ws[2][j] = a, b, a * b % p
                qs[3][j], qs[2][j] = 1, p - 1
                j += 1
            elif g == 1:                     # equality gate: (a-b)*eq + (a-b)*inv + eq - one
                a = rnd(); b = a if rng.random() < 0.5 else rnd()
                eq = 1 if a == b else 0
                inv = 0 if a == b else pow(a - b, -1, p)
                ws[0][j], ws[1][j], ws[2][j], ws[3][j], ws[4][j] = a, b, 1, eq, inv
                qs[8][j] = 1
                j += 1
            elif g == 2:                     # range-check accumulation: next.w0 = w0 + sum b_i r_i, then a free row
                acc = rnd()
                ws[0][j] = acc
                for i in range(15):
                    ws[1 + i][j] = rng.randrange(2)
                    rs[i][j] = pow(2, i, p)
                    acc = (acc + ws[1 + i][j] * rs[i][j]) % p
                qs[9][j] = 1
                ws[0][j + 1] = acc
                j += 2
            else:                            # five Poseidon rounds on a row, result on the next row's w0..w2
                st = [rnd(), rnd(), rnd()]
                ws[0][j], ws[1][j], ws[2][j] = st
                for rd in range(5):
                    rc = [rnd(), rnd(), rnd()]
                    rs[3 * rd][j], rs[3 * rd + 1][j], rs[3 * rd + 2][j] = rc
                    sb = [pow(x, 7, p) for x in st]
                    st = [(rc[i] + sum(M[3 * i + k] * sb[k] for k in range(3))) % p for i in range(3)]
                    if rd < 4:
                        ws[3 * rd + 3][j], ws[3 * rd + 4][j], ws[3 * rd + 5][j] = st
                ws[0][j + 1], ws[1][j + 1], ws[2][j + 1] = st
                qs[5][j] = 1
                j += 2
    else:
        raise ValueError(kind)
    cols = qs + ws + rs + ids + sig
    polys, evals = columns_to_polys(O, field, cols, lg_n)
    return dict(polys=polys, w_evals=evals[Q:Q + 3].copy(), public_inputs=O.ints_to_mont(field, pub).reshape(-1, 4),
                n_public=len(pub), lg_n=lg_n, field=field)
