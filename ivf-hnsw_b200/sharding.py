"""Host-side plumbing of the sharded search: which centroid range a rank owns and the
exchange of partial top-k lists.  Backend agnostic (`nccl` on the GPUs, `gloo` in the CPU tests); the compute --
partial scan and the merge -- stays in the CUDA library (search_partial_dev / merge_partials_dev)."""


def shard_range(rank, world, nc):
    """Rank `rank` of `world` owns the inverted lists of centroids [begin, end): equal contiguous ranges, the
    last ones shorter when world does not divide nc (a rank may own nothing)."""
    per = (nc + world - 1) // world
    begin = min(nc, rank * per)
    return begin, min(nc, begin + per)


def auto_list_shards(world, index_bytes, budget_bytes):
    """How many list shards a deployment of `world` GPUs needs: the smallest divisor of `world` whose shard of the
    inverted lists fits `budget_bytes` of one GPU's memory.  List sharding is the CAPACITY axis (every rank of a
    shard group must see every query of its group); the remaining factor world / list_shards is the THROUGHPUT
    axis: query groups that hold replicas of the lists and never talk to each other on the data path."""
    for s in range(1, world + 1):
        if world % s == 0 and index_bytes <= budget_bytes * s:
            return s
    return world


def grid_layout(rank, world, list_shards):
    """Rank -> (query_group, list_rank, ranks of its shard group) in a query_groups x list_shards grid; the ranks
    of one shard group are consecutive (same NVLink domain, and consecutive local ranks under torchrun)."""
    if list_shards < 1 or world % list_shards:
        raise ValueError("list_shards must divide the world size")
    qg = rank // list_shards
    return qg, rank % list_shards, list(range(qg * list_shards, (qg + 1) * list_shards))


def gather_partials(dist, part_d, part_seq, part_labels, out_d, out_seq, out_labels, group=None):
    """All-gather this rank's sorted partial top-k (distance, scan-order key, label), each [nq][k], into the
    [shards][nq][k] buffers every rank of the shard group then merges identically.  Three collectives of
    nq*k*(4+4+8) bytes."""
    for out, part in ((out_d, part_d), (out_seq, part_seq), (out_labels, part_labels)):
        # concatenated along dim 0 == [shards][nq][k] in memory (the flat form is what gloo accepts too)
        dist.all_gather_into_tensor(out.view(-1, *part.shape[1:]), part, group=group)


def query_slice(rank, world, nq):
    """Queries [begin, end) whose work tables rank `rank` prepares; `per` records per rank in the gathered buffer."""
    per = (nq + world - 1) // world
    begin = min(nq, rank * per)
    return begin, min(nq, begin + per), per


def gather_tables(dist, tables, rank, per, group=None):
    """In-place all-gather of the per-query work-table records: `tables` is [shards * per][record_bytes] and list
    rank r has filled rows [r * per, (r + 1) * per) (ivfhnsw_b200_prepare_dev); afterwards every rank of the shard
    group holds all records."""
    dist.all_gather_into_tensor(tables, tables[rank * per:(rank + 1) * per], group=group)


def exchange_list_meta(torch, dist, index, device, group=None):
    """Last step of a SHARDED index construction (ivfhnsw_b200_set_list_meta_dev): every rank has set_shard() its
    range, ingested only the points of its own lists and finalized; the per-list tables (sizes, sub-group sizes,
    alphas) are zero outside a rank's range, so one SUM all-reduce of each gives the whole tables, which every rank
    installs.  Afterwards pruning and the max_codes rule behave as on a rank that holds the complete index."""
    nc, nsubc = index.nc, getattr(index, "nsubc", 0)
    ls = torch.zeros(nc, dtype=torch.int32, device=device)
    ss = torch.zeros(max(nc * nsubc, 1), dtype=torch.int32, device=device)
    al = torch.zeros(nc, dtype=torch.float32, device=device)
    index.get_list_meta_dev(ls, ss if nsubc else None, al if nsubc else None)
    index.sync()
    dist.all_reduce(ls, group=group)
    if nsubc:
        dist.all_reduce(ss, group=group)
        dist.all_reduce(al, group=group)
    index.set_list_meta_dev(ls, ss if nsubc else None, al if nsubc else None)
    return int(ls.sum().item())


class PackedShardedSearch:
    """The list-sharded search of SURVEY.md 8(e) without an all-gather of whole records.

    Every rank is the HOME of `nq_home` queries per step (rows [rank * nq_home, (rank + 1) * nq_home) of the batch
    `q_all`, which every rank holds).  Per step:
      1. home: coarse stage + pruning pass + max_codes cut of the home queries      (ivfhnsw_b200_prepare_dev)
      2. home: records -> one packed stream per owning rank                          (ivfhnsw_b200_pack_records_dev)
      3. the stream lengths are all-gathered (world x world numbers; the ONE host read of the step), then ONE
         all-to-all-v moves the streams and one fixed all-to-all the per-query (count, offset) tables
      4. every rank scans the lists it owns for the queries of all ranks            (ivfhnsw_b200_scan_packed_dev)
      5. the partial top-k lists go back to the queries' homes (fixed all-to-all) and are merged there by
         (distance, global scan order): bit-identical to a single-GPU search        (ivfhnsw_b200_merge_partials_dev)
    Backend agnostic (`nccl` on the GPUs, `gloo` in the CPU tests of the exchange arithmetic)."""

    def __init__(self, torch, dist, index, rank, world, nq_home, k, device, group=None):
        self.torch, self.dist, self.idx, self.rank, self.world = torch, dist, index, rank, world
        self.nq_home, self.k, self.group = nq_home, k, group
        nqt = nq_home * world
        i32, f32, i64 = torch.int32, torch.float32, torch.int64
        rec_words = index.tables_bytes() // 4
        self.tables = torch.empty(nq_home, rec_words, dtype=i32, device=device)
        self.packed = torch.empty(nq_home * rec_words, dtype=i32, device=device)
        self.cnt = torch.empty(world, nq_home, dtype=i32, device=device)            # [part][query] probes
        self.off = torch.empty(world, nq_home, dtype=i32, device=device)            # [part][query] word offset in its stream
        self.meta_in = torch.empty(world, 2, nq_home, dtype=i32, device=device)     # [source][cnt | off][query]
        self.pwords = torch.empty(world, dtype=i64, device=device)
        self.pw_all = torch.empty(world, world, dtype=i64, device=device)
        self.pD = torch.empty(nqt, k, dtype=f32, device=device)
        self.pS = torch.empty(nqt, k, dtype=i32, device=device)
        self.pL = torch.empty(nqt, k, dtype=i64, device=device)
        self.gD = torch.empty(world, nq_home, k, dtype=f32, device=device)
        self.gS = torch.empty(world, nq_home, k, dtype=i32, device=device)
        self.gL = torch.empty(world, nq_home, k, dtype=i64, device=device)
        self.D = torch.empty(nq_home, k, dtype=f32, device=device)
        self.L = torch.empty(nq_home, k, dtype=i64, device=device)
        self.last_recv_words = 0

    @staticmethod
    def recv_layout(torch, meta_in, recv_words):
        """(qoff, qnp) of all queries from the received (count, offset) tables: source s's stream starts where the
        streams of the sources before it end."""
        base = torch.cumsum(recv_words, 0) - recv_words                     # exclusive scan, int64
        qoff = (meta_in[:, 1, :].to(torch.int64) + base[:, None]).to(torch.int32)
        return qoff.reshape(-1).contiguous(), meta_in[:, 0, :].reshape(-1).contiguous()

    def search(self, q_all):
        torch, dist, idx, r, W, nh = self.torch, self.dist, self.idx, self.rank, self.world, self.nq_home
        q_home = q_all[r * nh:(r + 1) * nh]
        idx.prepare_dev(q_home, 0, nh, self.tables)
        idx.pack_records_dev(nh, self.tables, W, self.packed, self.cnt, self.off, self.pwords)
        dist.all_gather_into_tensor(self.pw_all.view(-1), self.pwords, group=self.group)
        pw = self.pw_all.cpu()                                              # the step's one host read
        send = [int(v) for v in pw[r]]
        recv = [int(v) for v in pw[:, r]]
        self.last_recv_words = sum(recv)
        blob = torch.empty(max(1, sum(recv)), dtype=torch.int32, device=q_all.device)
        dist.all_to_all_single(blob[:sum(recv)], self.packed[:sum(send)], recv, send, group=self.group)
        dist.all_to_all_single(self.meta_in, torch.stack((self.cnt, self.off), 1), group=self.group)
        qoff, qnp = self.recv_layout(torch, self.meta_in, self.pw_all[:, r])
        idx.scan_packed_dev(self.k, q_all, blob, qoff, qnp, self.pD, self.pS, self.pL)
        for out, part in ((self.gD, self.pD), (self.gS, self.pS), (self.gL, self.pL)):
            dist.all_to_all_single(out.view(-1, self.k), part, group=self.group)
        idx.merge_partials_dev(self.k, nh, W, self.gD, self.gS, self.gL, self.D, self.L)
        return self.D, self.L
