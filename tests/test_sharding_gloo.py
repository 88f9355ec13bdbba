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


# ------------------------------------------------------------------ PackedShardedSearch (all-to-all-v protocol) on gloo
class _FakeIndex:
    """Stands in for the CUDA library in the world-2 CPU test of sharding.PackedShardedSearch: records are synthetic
    (IVFADC layout, no sub-groups), the pack step is the numpy statement of ivfhnsw_b200_pack_records_dev
    (tests/pack_ref.py), a 'scan' turns every received probe into one candidate."""
    NPROBE, NC = 6, 40

    def __init__(self, rank, world):
        self.rank, self.world = rank, world

    def tables_bytes(self):
        import pack_ref
        return pack_ref.record_words(self.NPROBE, 0) * 4

    @classmethod
    def records_of(cls, q_home):
        import pack_ref
        P = cls.NPROBE
        rec = np.zeros((q_home.shape[0], pack_ref.record_words(P, 0)), np.uint32)
        for i, q in enumerate(q_home[:, 0].numpy().astype(np.int64)):
            rng = np.random.default_rng(int(q))
            cid = rng.permutation(cls.NC)[:P].astype(np.uint32)
            n = rng.integers(0, 3, P).astype(np.uint32)              # 0 = skipped by the max_codes rule
            rec[i, 0:P] = cid * 10                                    # slot
            rec[i, P:2 * P] = n
            rec[i, 2 * P:3 * P] = (np.float32(q) + cid.astype(np.float32)).view(np.uint32)   # base = the candidate's distance
            rec[i, 3 * P:4 * P] = np.arange(P, dtype=np.uint32) * 100   # seq
            rec[i, 4 * P:5 * P] = cid
        return rec

    def prepare_dev(self, q_home, lo, hi, tables):
        tables.copy_(torch.from_numpy(self.records_of(q_home).view(np.int32)))

    def pack_records_dev(self, nq, tables, nparts, packed, cnt, off, pwords):
        import pack_ref
        streams, c, o, pw = pack_ref.pack_records(tables.numpy().view(np.uint32), self.NPROBE, 0, self.NC, nparts)
        flat = np.concatenate(streams)
        packed[:flat.size] = torch.from_numpy(flat.view(np.int32))
        cnt.copy_(torch.from_numpy(c.view(np.int32))); off.copy_(torch.from_numpy(o.view(np.int32)))
        pwords.copy_(torch.from_numpy(pw.astype(np.int64)))

    def scan_packed_dev(self, k, q_all, blob, qoff, qnp, pD, pS, pL):
        sharding = importlib.import_module("ivf-hnsw_b200.sharding")
        b, e = sharding.shard_range(self.rank, self.world, self.NC)
        pD.fill_(np.finfo(np.float32).max); pL.fill_(-1)
        pS.copy_(torch.arange(k, dtype=torch.int32).flip(0).expand_as(pS) * -1 + np.iinfo(np.int32).max)
        blob_np = blob.numpy().view(np.uint32)
        for q in range(q_all.shape[0]):
            npq, o = int(qnp[q]), int(qoff[q])
            r = blob_np[o:o + 5 * npq]
            cid = r[4 * npq:5 * npq]
            assert np.all((cid >= b) & (cid < e))                     # only probes this rank owns arrive here
            cand = sorted(zip(r[2 * npq:3 * npq].view(np.float32).tolist(), r[3 * npq:4 * npq].tolist(), cid.tolist()))[:k]
            for j, (dd, ss, cc) in enumerate(cand):
                pD[q, j], pS[q, j], pL[q, j] = dd, ss, cc

    def merge_partials_dev(self, k, nq, nparts, gD, gS, gL, D, L):
        d, l = merge_host(gD.numpy(), gS.numpy(), gL.numpy(), k)
        D.copy_(torch.from_numpy(d)); L.copy_(torch.from_numpy(l))


def _packed_worker(rank, world, port_no, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    nq_home, k = 7, 4
    q_all = torch.arange(1, world * nq_home + 1, dtype=torch.float32)[:, None].repeat(1, 3) * 1000.0
    fake = _FakeIndex(rank, world)
    ps = sharding.PackedShardedSearch(torch, dist, fake, rank, world, nq_home, k, torch.device("cpu"))
    D, L = ps.search(q_all)
    # expected: all scanned probes of the home queries, k best by (distance, seq)
    rec = _FakeIndex.records_of(q_all[rank * nq_home:(rank + 1) * nq_home])
    P = _FakeIndex.NPROBE
    ok = True
    for i in range(nq_home):
        keep = rec[i, P:2 * P] != 0
        cand = sorted(zip(rec[i, 2 * P:3 * P].view(np.float32)[keep].tolist(), rec[i, 3 * P:4 * P][keep].tolist(),
                          rec[i, 4 * P:5 * P][keep].tolist()))[:k]
        want_l = [c for _, _, c in cand] + [-1] * (k - len(cand))
        ok = ok and L[i].tolist() == want_l and np.allclose(D[i].numpy()[:len(cand)], [c[0] for c in cand])
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_packed_sharded_search_protocol_on_gloo():
    """World size 2: every rank prepares its own queries, the packed per-owner streams travel by all-to-all-v, every
    rank 'scans' only what it owns, the partial lists return home and merge to the answer over ALL probes."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port_no = s.getsockname()[1]
    world = 2
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_packed_worker, args=(world, port_no, out), nprocs=world, join=True)
        assert all(out.get(r) for r in range(world)), dict(out)


class _FakeShardIndex:
    """Per-list tables of a rank that ingested only the lists it owns (zero elsewhere), for the world-2 CPU test of
    sharding.exchange_list_meta."""
    nc, nsubc = 10, 4

    def __init__(self, rank, world):
        sharding = importlib.import_module("ivf-hnsw_b200.sharding")
        self.lo, self.hi = sharding.shard_range(rank, world, self.nc)
        self.installed = None

    @classmethod
    def whole(cls):
        sub = (np.arange(cls.nc * cls.nsubc, dtype=np.int32).reshape(cls.nc, cls.nsubc) * 7) % 5
        return sub.sum(1).astype(np.int32), sub.reshape(-1), (np.arange(cls.nc, dtype=np.float32) + 1) / 16

    def get_list_meta_dev(self, ls, ss, al):
        wl, ws, wa = self.whole()
        own = np.zeros(self.nc, bool); own[self.lo:self.hi] = True
        ls.copy_(torch.from_numpy(np.where(own, wl, 0).astype(np.int32)))
        ss.copy_(torch.from_numpy(np.where(np.repeat(own, self.nsubc), ws, 0).astype(np.int32)))
        al.copy_(torch.from_numpy(np.where(own, wa, 0).astype(np.float32)))

    def set_list_meta_dev(self, ls, ss, al):
        self.installed = (ls.numpy().copy(), ss.numpy().copy(), al.numpy().copy())

    def sync(self):
        pass


def _meta_worker(rank, world, port_no, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    fake = _FakeShardIndex(rank, world)
    total = sharding.exchange_list_meta(torch, dist, fake, torch.device("cpu"))
    wl, ws, wa = _FakeShardIndex.whole()
    out[rank] = bool(total == int(wl.sum()) and np.array_equal(fake.installed[0], wl) and
                     np.array_equal(fake.installed[1], ws) and np.array_equal(fake.installed[2], wa))
    dist.destroy_process_group()


def test_sharded_construction_metadata_exchange_on_gloo():
    """World size 2: the per-list tables of ranks that ingested disjoint list ranges sum to the whole tables, and
    every rank installs the same whole tables (the last step of a sharded index construction)."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port_no = s.getsockname()[1]
    world = 2
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_meta_worker, args=(world, port_no, out), nprocs=world, join=True)
        assert all(out.get(r) for r in range(world)), dict(out)
