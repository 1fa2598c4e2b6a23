"""CPU tests of the N>1 host logic with the gloo backend, world_size 2 (and in-process checks)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from ransaclib_b200 import sharding


def test_hypothesis_ranges_cover_and_ascend():
    for k, w in [(65536, 8), (10, 4), (7, 8), (1, 2)]:
        spans = [sharding.hypothesis_range(k, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == k
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    assert sorted(sum((sharding.query_indices(10, r, 4) for r in range(4)), [])) == list(range(10))


def test_select_winner_rules():
    g = torch.tensor([[5.0, 7, 0], [3.0, 40, 1], [3.0, 12, 2], [float("nan"), 1, 3], [sharding.DBL_MAX, -1, 4]],
                     dtype=torch.float64)
    w, ok = sharding.select_winner(g)
    assert int(w) == 2 and bool(ok)  # lowest score, then lowest id; NaN and DBL_MAX never win
    w, ok = sharding.select_winner(g[3:])
    assert not bool(ok)


def _worker(rank, world, port, scores, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist

    dist.init_process_group("gloo", rank=rank, world_size=world)
    k = len(scores)
    lo, hi = sharding.hypothesis_range(k, rank, world)
    local = torch.tensor(scores[lo:hi], dtype=torch.float64)
    usable = torch.where(local < sharding.DBL_MAX, local, torch.full_like(local, float("inf")))
    lid = torch.argmin(usable) if len(local) else torch.tensor(-1)
    has = bool((local < sharding.DBL_MAX).any())
    lid = lid if has else torch.tensor(-1)
    model = torch.full((3,), float(rank), dtype=torch.float64)
    rec = sharding.pack_record(local[lid.clamp(min=0)], lid, lo, model)
    # gloo has no all_gather_into_tensor on every build: use the list form here
    outs = [torch.empty_like(rec) for _ in range(world)]
    dist.all_gather(outs, rec)
    g = torch.stack(outs)
    w, ok = sharding.select_winner(g)
    q.put((rank, int(g[w, 1].item()), bool(ok), float(g[w, 2].item())))
    dist.destroy_process_group()


@pytest.mark.parametrize("scores", [
    [9.0, 4.0, 7.0, 4.0, 8.0, 4.0],                       # tie across ranks: lowest global id wins
    [sharding.DBL_MAX, sharding.DBL_MAX, 2.0, 1.0, 1.0, 3.0],  # rank 0 has no valid hypothesis
    [float("nan"), 5.0, 6.0, float("nan"), 5.0, 7.0],
])
def test_cross_rank_argmin_gloo_world2(scores):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, scores, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    arr = np.array(scores)
    arr = np.where(np.isnan(arr) | (arr >= sharding.DBL_MAX), np.inf, arr)
    want = int(np.argmin(arr))
    for rank, gid, ok, owner in res:
        assert ok and gid == want
        assert owner == float(0 if want < 3 else 1)  # the model came from the owning rank
