"""World-size-2 gloo tests (CPU) of the multi-GPU host logic in halo_b200/sharded.py: range partition, partial-sum
exchange and combine, round-robin batches.  The per-rank compute is injected (the oracle stands in for the GPU), so
what is under test is exactly the part that does not need a device."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here); sys.path.insert(0, os.path.dirname(here))
    import torch.distributed as dist
    import oracle_lib as O
    from halo_b200 import sharded
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        O.set_threads(2)
        k = 1001                                               # ragged: ranges of 501 and 500 points
        pts = O.points_arith(0, k)
        sc = O.random_scalars(0, k, 5)
        combine = lambda c, parts: __import__("functools").reduce(lambda a, b: O.point_add(c, a, b), list(parts))
        got = sharded.msm_sharded(dist, 0, sc, pts, local_msm=lambda c, s, b, off: O.msm(c, b, s), combine=combine)
        exp = O.msm(0, pts, sc)
        ok_msm = bool((O.point_to_affine(0, got.reshape(1, 12)) == O.point_to_affine(0, exp.reshape(1, 12))).all())
        # batch of 5 commit columns over shared bases, round-robin
        cols = np.stack([O.random_scalars(0, 64, 10 + i) for i in range(5)])
        res = sharded.msm_batch_sharded(dist, 0, cols, local_batch=lambda c, cs, off: np.stack([O.msm(c, pts[:64], x) for x in cs]))
        ok_batch = all((O.point_to_affine(0, res[i].reshape(1, 12)) == O.point_to_affine(0, O.msm(0, pts[:64], cols[i]).reshape(1, 12))).all()
                       for i in range(5))
        polys = np.stack([O.random_scalars(0, 32, 20 + i) for i in range(3)])
        tr = sharded.ntt_batch_sharded(dist, 0, polys, local_ntt=lambda f, x, inv: O.ntt(f, x, inverse=inv))
        ok_ntt = all((tr[i] == O.ntt(0, polys[i])).all() for i in range(3))
        q.put((rank, ok_msm, ok_batch, ok_ntt, sharded.point_range(k, rank, world)))
    except Exception as e:                                     # surface the failure instead of a queue timeout
        q.put((rank, False, False, False, repr(e)))
        raise
    finally:
        dist.destroy_process_group()


def test_point_range_partition():
    from halo_b200 import sharded
    for k in (0, 1, 7, 8, 1000, 1 << 20):
        for world in (1, 2, 3, 8):
            r = [sharded.point_range(k, g, world) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == k
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in r) - min(h - l for l, h in r) <= 1
            assert sorted(sum((sharded.items_of(k % 50, g, world) for g in range(world)), [])) == list(range(k % 50))


@pytest.mark.timeout(300)
def test_sharded_paths_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[0] for r in res] == [0, 1]
    assert all(r[1] and r[2] and r[3] for r in res), res
    assert res[0][4] == (0, 501) and res[1][4] == (501, 1001)
