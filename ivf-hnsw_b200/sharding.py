"""Host-side plumbing of the sharded search: which centroid range a rank owns and the
exchange of partial top-k lists.  Backend agnostic (`nccl` on the GPUs, `gloo` in the CPU tests); the compute --
partial scan and the merge -- stays in the CUDA library (search_partial_dev / merge_partials_dev)."""


def shard_range(rank, world, nc):
    """Rank `rank` of `world` owns the inverted lists of centroids [begin, end): equal contiguous ranges, the
    last ones shorter when world does not divide nc (a rank may own nothing)."""
    per = (nc + world - 1) // world
    begin = min(nc, rank * per)
    return begin, min(nc, begin + per)


def gather_partials(dist, part_d, part_seq, part_labels, out_d, out_seq, out_labels):
    """All-gather this rank's sorted partial top-k (distance, scan-order key, label), each [nq][k], into the
    [world][nq][k] buffers every rank then merges identically.  Three collectives of nq*k*(4+4+8) bytes."""
    for out, part in ((out_d, part_d), (out_seq, part_seq), (out_labels, part_labels)):
        # concatenated along dim 0 == [world][nq][k] in memory (the flat form is what gloo accepts too)
        dist.all_gather_into_tensor(out.view(-1, *part.shape[1:]), part)


def query_slice(rank, world, nq):
    """Queries [begin, end) whose work tables rank `rank` prepares; `per` records per rank in the gathered buffer."""
    per = (nq + world - 1) // world
    begin = min(nq, rank * per)
    return begin, min(nq, begin + per), per


def gather_tables(dist, tables, rank, per):
    """In-place all-gather of the per-query work-table records: `tables` is [world * per][record_bytes] and rank r
    has filled rows [r * per, (r + 1) * per) (ivfhnsw_b200_prepare_dev); afterwards every rank holds all records."""
    dist.all_gather_into_tensor(tables, tables[rank * per:(rank + 1) * per])
