"""Multi-GPU partitioning of the ORB front-end hot path (one process per GPU, torch.distributed for the plumbing).

Only two things on this path shard:
  * extraction: independent frames -> frame i goes to rank i mod G, no collective on the data path;
  * brute-force top-2 over a keyframe-descriptor database: rows are block-sharded [g*M/G, (g+1)*M/G), every rank scans
    its shard (orbm_bf_top2_device) and ONE all-gather of the per-shard (d1, idx1, d2) triples is followed by the same
    on-device merge on every rank (orbm_merge_top2_device).
Guided matching (SearchByProjection etc.) stays on the GPU that holds the frame: replicas only.

This module holds the host-side logic (who owns what, the gather layout); it computes nothing itself.
"""
import torch
import torch.distributed as dist


def shard_rows(total_rows, world, rank):
    """Contiguous ascending row range [begin, end) of shard `rank`; earlier shards take the remainder rows."""
    if world < 1 or not (0 <= rank < world) or total_rows < 0:
        raise ValueError("bad shard request")
    base, rem = divmod(total_rows, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def frames_for_rank(nframes, world, rank):
    """Frame indices handled by `rank`: round-robin, frame i -> rank i mod world (stereo pairs stay together by passing
    the number of PAIRS)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank")
    return list(range(rank, nframes, world))


def gather_top2(d1, idx1, d2, group=None):
    """All-gather the per-shard top-2 triples.  Inputs: (nq,) int32 / int64 / int32 tensors on any device the process
    group supports.  Returns ([G, nq] d1, [G, nq] idx1, [G, nq] d2) laid out by ascending rank = ascending row range,
    which is the layout orbm_merge_top2_device expects.  One collective of 16 bytes per query."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    nq = d1.shape[0]
    if world == 1:
        return d1.view(1, nq), idx1.view(1, nq), d2.view(1, nq)
    # pack into one int64 buffer so the exchange is a single all-gather: [d1 | d2] in one word, idx1 in the other
    packed = torch.empty((nq, 2), dtype=torch.int64, device=d1.device)
    packed[:, 0] = (d1.to(torch.int64) << 32) | d2.to(torch.int64)
    packed[:, 1] = idx1
    out = torch.empty((world, nq, 2), dtype=torch.int64, device=d1.device)
    dist.all_gather_into_tensor(out.view(world * nq, 2), packed, group=group)
    d1_all = (out[:, :, 0] >> 32).to(torch.int32).contiguous()
    d2_all = (out[:, :, 0] & 0xFFFFFFFF).to(torch.int32).contiguous()
    return d1_all, out[:, :, 1].contiguous(), d2_all


class ShardedDescriptorDB:
    """Row-sharded keyframe-descriptor database: rank g holds rows shard_rows(total, G, g) in HBM.

    top2(q_dev) -> (d1, idx1, d2) device tensors holding the GLOBAL sequential-scan result on every rank."""

    def __init__(self, total_rows, device_index, group=None):
        import orbb200
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.begin, self.end = shard_rows(total_rows, self.world, self.rank)
        self.dev = torch.device("cuda", device_index)
        self.db = orbb200.DescriptorDB(max(self.end - self.begin, 1), row_offset=self.begin, device=device_index)
        self._orbb = orbb200

    def append_device(self, rows_dev):
        assert rows_dev.dtype == torch.uint8 and rows_dev.is_cuda and rows_dev.is_contiguous()
        self.db.append_device(rows_dev.data_ptr(), rows_dev.shape[0])

    def top2(self, q_dev):
        nq = q_dev.shape[0]
        d1 = torch.empty(nq, dtype=torch.int32, device=self.dev)
        i1 = torch.empty(nq, dtype=torch.int64, device=self.dev)
        d2 = torch.empty(nq, dtype=torch.int32, device=self.dev)
        st = torch.cuda.current_stream(self.dev).cuda_stream
        self.db.top2_device(q_dev.data_ptr(), nq, d1.data_ptr(), i1.data_ptr(), d2.data_ptr(), st)
        if self.world == 1:
            return d1, i1, d2
        d1a, i1a, d2a = gather_top2(d1, i1, d2, self.group)
        self._orbb.merge_top2_device(d1a.data_ptr(), i1a.data_ptr(), d2a.data_ptr(), self.world, nq,
                                     d1.data_ptr(), i1.data_ptr(), d2.data_ptr(), st)
        return d1, i1, d2
