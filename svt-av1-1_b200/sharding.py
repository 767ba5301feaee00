"""Host-side work partitioning for N GPUs (one process per GPU). ME reads only source pictures, so the path
shards without any data-path collective:

  stream_shard    N independent picture streams (BASELINE configs[4]): whole streams per rank
  picture_shard   one stream, mini-GOPs dealt to ranks; every rank uploads the pictures its jobs read
                  (references are replicated, the reference's PA pictures are read-only)
  picture_shard_contiguous   one stream, contiguous runs of the decode order per rank, with the owner of every picture
  sb_row_shard    one large picture (4K), contiguous SB rows per rank; results are disjoint row ranges
  sb_shard        the same by SBs in raster order (balanced to one SB)
  aggregate       whole-job throughput = units of all ranks / max-over-ranks device time

Only `aggregate` communicates (two tiny all-reduces outside the timed region); it works on any
torch.distributed backend (NCCL on the GPU box, gloo in the CPU tests).
"""


def stream_shard(n_streams, world, rank):
    """Stream ids owned by `rank` (contiguous blocks, sizes differ by at most one)."""
    base, extra = divmod(n_streams, world)
    first = rank * base + min(rank, extra)
    return list(range(first, first + base + (1 if rank < extra else 0)))


def picture_shard(jobs, world, rank, gop=8):
    """jobs: [(poc, temporal_layer, ref0_poc, ref1_poc)] in decode order. Mini-GOP g goes to rank g % world.
    Returns (jobs of this rank, sorted POCs this rank must hold in HBM)."""
    mine = [j for j in jobs if ((j[0] - 1) // gop) % world == rank]
    need = sorted({p for j in mine for p in (j[0], j[2], j[3])})
    return mine, need


def picture_shard_contiguous(jobs, world, rank):
    """One stream, pictures dealt to ranks as contiguous runs of the decode order (sizes differ by at most one picture).
    Returns (jobs of this rank, sorted POCs this rank must hold, {poc: owner rank}): the owner of a picture is the rank
    that motion-estimates it (POC 0, the intra picture, belongs to rank 0); the owner uploads it from the host and the
    other ranks that reference it import it from the owner's HBM (svtme_picture_upload_ipc)."""
    def span(r):
        base, extra = divmod(len(jobs), world)
        first = r * base + min(r, extra)
        return first, first + base + (1 if r < extra else 0)
    owner = {0: 0}
    for r in range(world):
        lo, hi = span(r)
        for j in jobs[lo:hi]:
            owner[j[0]] = r
    lo, hi = span(rank)
    mine = jobs[lo:hi]
    need = sorted({p for j in mine for p in (j[0], j[2], j[3])})
    return mine, need, owner


def sb_shard(n_sb, world, rank):
    """(first_sb, n_sbs) of the contiguous raster range of SBs owned by `rank` (SvtMeJob.range_in_sbs = 1)."""
    base, extra = divmod(n_sb, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def sb_row_shard(sb_rows, world, rank):
    """(first_row, n_rows) of the contiguous SB-row range owned by `rank`."""
    base, extra = divmod(sb_rows, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def aggregate(units_local, ms_local, dist=None, device="cpu"):
    """(total units, max ms, units per second) over all ranks; dist = torch.distributed or None for one rank."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return units_local, ms_local, units_local / (ms_local / 1e3)
    import torch
    t = torch.tensor([float(ms_local)], dtype=torch.float64, device=device)
    u = torch.tensor([float(units_local)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return u.item(), t.item(), u.item() / (t.item() / 1e3)
