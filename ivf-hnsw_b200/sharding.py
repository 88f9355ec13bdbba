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
