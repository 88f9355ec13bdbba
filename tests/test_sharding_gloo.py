"""The N > 1 path on CPU: two gloo processes run the sharding protocol of SURVEY.md 8(e) -- lists owned by
centroid range, partial top-k per shard, all-gather into [world][nq][k], identical merge by (distance, scan order)
on every rank -- with the C port standing in for the per-shard scan.  The merged answer must equal the single-shard
answer of the port.  (The CUDA merge kernel itself is covered by the gpu tests with scan_split > 1 and by
bench.py --gpus N.)"""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import load_golden

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))


def merge_host(gD, gS, gL, k):
    """[world][nq][k] -> [nq][k] by (distance, scan-order key): what merge_partials_dev computes."""
    world, nq, _ = gD.shape
    D = np.empty((nq, k), np.float32)
    L = np.empty((nq, k), np.int64)
    for q in range(nq):
        d = gD[:, q].ravel(); s = gS[:, q].ravel(); lab = gL[:, q].ravel()
        part = np.repeat(np.arange(world), gD.shape[2])
        o = np.lexsort((part, s, d))[:k]
        D[q], L[q] = d[o], lab[o]
    return D, L


def _worker(rank, world, port_no, name, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    import oracle as O
    g, ix, runs = load_golden(name)
    k, nprobe = int(g["k"]), int(g["nprobe"])
    tag, mc, pr = runs[0]
    port = O.Port()
    # the port scans everything the reference rule reaches: K = all codes gives the full ordered candidate list
    K = int(min(ix["ids"].shape[0], 4096))
    Df, Lf = port.search(ix, g["Q"], g["cd_" + tag], g["ci_" + tag], K, nprobe, mc, pr)
    o = np.argsort(Df, axis=1, kind="stable")
    Df, Lf = np.take_along_axis(Df, o, 1), np.take_along_axis(Lf, o, 1)
    # owner of every base vector = the list it sits in
    owner_list = np.repeat(np.arange(ix["nc"]), np.diff(ix["list_off"]))
    list_of_id = np.empty(int(ix["ids"].max()) + 1, np.int64)
    list_of_id[ix["ids"]] = owner_list
    begin, end = sharding.shard_range(rank, world, int(ix["nc"]))
    nq = Df.shape[0]
    pD = np.full((nq, k), np.finfo(np.float32).max, np.float32)
    pS = np.empty((nq, k), np.int32); pL = np.full((nq, k), -1, np.int64)
    for q in range(nq):
        valid = Lf[q] >= 0
        mine = valid & (list_of_id[np.where(valid, Lf[q], 0)] >= begin) & (list_of_id[np.where(valid, Lf[q], 0)] < end)
        pos = np.nonzero(mine)[0][:k]
        pD[q, :pos.size], pL[q, :pos.size] = Df[q, pos], Lf[q, pos]
        pS[q, :pos.size] = pos                                   # global scan-order stand-in: rank in the full list
        pS[q, pos.size:] = np.iinfo(np.int32).max - np.arange(k - pos.size)[::-1]
    tD, tS, tL = torch.from_numpy(pD), torch.from_numpy(pS), torch.from_numpy(pL)
    gD = torch.empty(world, nq, k); gS = torch.empty(world, nq, k, dtype=torch.int32)
    gL = torch.empty(world, nq, k, dtype=torch.int64)
    sharding.gather_partials(dist, tD, tS, tL, gD, gS, gL)
    D, L = merge_host(gD.numpy(), gS.numpy(), gL.numpy(), k)
    ok = bool(np.array_equal(L, Lf[:, :k]) and np.array_equal(D, Df[:, :k]))
    # every rank must hold the same merged answer
    chk = torch.from_numpy(L.copy())
    dist.broadcast(chk, src=0)
    ok = ok and bool(np.array_equal(chk.numpy(), L))
    out[rank] = ok
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["adc_d32", "grp_d32"])
def test_sharded_partials_merge_to_the_single_shard_answer(name):
    import oracle as O
    O.build()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port_no = s.getsockname()[1]
    world = 2
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, port_no, name, out), nprocs=world, join=True)
        assert all(out.get(r) for r in range(world)), dict(out)


def test_shard_ranges_cover_all_centroids():
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    for nc in (1, 7, 256, 65536, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            r = [sharding.shard_range(i, world, nc) for i in range(world)]
            assert r[0][0] == 0 and r[-1][1] == nc
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert all(0 <= a <= b <= nc for a, b in r)


def test_grid_layout_and_capacity_rule():
    """query_groups x list_shards grid: shard groups are consecutive ranks; the capacity rule picks the smallest
    divisor of the world whose list shard fits the per-GPU budget."""
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    assert sharding.grid_layout(5, 8, 2) == (2, 1, [4, 5])
    assert sharding.grid_layout(3, 8, 8) == (0, 3, list(range(8)))
    assert sharding.grid_layout(3, 8, 1) == (3, 0, [3])
    with pytest.raises(ValueError):
        sharding.grid_layout(0, 8, 3)
    gb = 1 << 30
    assert sharding.auto_list_shards(8, 22 * gb, 100 * gb) == 1        # 1B x 16-byte codes: replicas
    assert sharding.auto_list_shards(8, 250 * gb, 100 * gb) == 4
    assert sharding.auto_list_shards(6, 250 * gb, 100 * gb) == 3
    assert sharding.auto_list_shards(8, 5000 * gb, 100 * gb) == 8
    # every centroid is owned exactly once inside a shard group
    for R in (1, 2, 3, 8):
        cover = np.zeros(1000, np.int32)
        for r in range(R):
            b, e = sharding.shard_range(r, R, 1000)
            cover[b:e] += 1
        assert (cover == 1).all()
