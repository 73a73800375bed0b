"""Multi-GPU sharding of the scan: contiguous template segments + hit gather.

One process per GPU (torchrun), `torch.distributed` for the plumbing.  The scan
has no data-path collective: a template is cut into G contiguous segments of
window starts [b_g, e_g); rank g uploads bases [b_g, e_g + L_max - 1) — a
(p_len-1) halo — and scans it alone.  The only exchange is at the
end: hit records (16 B each) are all-gathered with a variable-length gather
(counts first, then padded payload) over NCCL on GPUs, gloo on CPU.
"""
import numpy as np

from . import _native as nat


def segments(t_len, world_size):
    """Window-start ranges [b, e) per rank; multiples of 32 so that every rank's
    planes stay word aligned."""
    words = (t_len + 31) // 32
    per = (words + world_size - 1) // world_size
    out = []
    for g in range(world_size):
        b = min(t_len, g * per * 32)
        e = min(t_len, (g + 1) * per * 32)
        out.append((b, e))
    return out


def shard_bounds(t_len, world_size, rank, max_primer_len):
    """(base_begin, base_end, start_begin, start_end) of this rank's slice."""
    b, e = segments(t_len, world_size)[rank]
    return b, min(t_len, e + max_primer_len - 1), b, e


def localize_hits(hits, shard_begin, shard_len, t_len, primer_lens):
    """Translate hits found on a shard (scanned as its own record) into
    whole-template coordinates.  Forward: pos + shard_begin.  Reverse strand
    positions index reverse_complement(), so they are mirrored:
    pos_global = t_len - L - (shard_begin + (shard_len - L - pos_local))."""
    out = hits.copy()
    L = np.asarray(primer_lens, dtype=np.int64)[hits['primer']]
    pos = hits['pos'].astype(np.int64)
    fwd_start = np.where(hits['strand'] == 0, pos, shard_len - L - pos) + shard_begin
    out['pos'] = np.where(hits['strand'] == 0, fwd_start, t_len - L - fwd_start).astype(np.uint64)
    return out


def keep_owned(hits, start_begin, start_end, t_len, primer_lens):
    """Drop hits whose forward window start lies in the halo (owned by the next rank)."""
    L = np.asarray(primer_lens, dtype=np.int64)[hits['primer']]
    pos = hits['pos'].astype(np.int64)
    fwd_start = np.where(hits['strand'] == 0, pos, t_len - L - pos)
    return hits[(fwd_start >= start_begin) & (fwd_start < start_end)]


def all_gather_records(arr, group=None, device=None):
    """Variable-length all-gather of a structured numpy array (allgatherv).

    NCCL has no allgatherv: the per-rank counts are gathered first, the payload
    is padded to the maximum count and gathered with one all_gather_into_tensor,
    then trimmed.  Payloads are KB-MB: latency-, not bandwidth-bound."""
    import torch
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return arr.copy()
    world = dist.get_world_size(group)
    backend = dist.get_backend(group)
    if device is None:
        device = torch.device('cuda', torch.cuda.current_device()) if backend == 'nccl' else torch.device('cpu')
    itemsize = arr.dtype.itemsize
    count = torch.tensor([len(arr)], dtype=torch.int64, device=device)
    counts = torch.zeros(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(counts, count, group=group)
    counts = counts.cpu().numpy()
    m = int(counts.max())
    if m == 0:
        return arr[:0].copy()
    raw = np.zeros(m * itemsize, dtype=np.uint8)
    raw[:len(arr) * itemsize] = np.frombuffer(arr.tobytes(), dtype=np.uint8)
    send = torch.from_numpy(raw).to(device)
    recv = torch.empty(world * m * itemsize, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(recv, send, group=group)
    recv = recv.cpu().numpy()
    parts = []
    for g in range(world):
        chunk = recv[g * m * itemsize:(g * m + int(counts[g])) * itemsize]
        parts.append(np.frombuffer(chunk.tobytes(), dtype=arr.dtype))
    return np.concatenate(parts)


def sort_hits(hits):
    """The order dp_scan uses: (primer, strand, record, pos)."""
    return np.sort(hits, order=['primer', 'strand', 'record', 'pos'])


class ShardedTemplate(object):
    """This rank's slice of one long template, resident on this rank's GPU."""

    def __init__(self, ctx, text, rank, world_size, max_primer_len=nat.DP_MAX_PRIMER_LEN):
        """`text`: the whole template (str / bytes / uint8 array); only this
        rank's slice is uploaded."""
        self.t_len = len(text)
        self.rank, self.world_size = rank, world_size
        self.base_begin, self.base_end, self.start_begin, self.start_end = \
            shard_bounds(self.t_len, world_size, rank, max_primer_len)
        piece = text[self.base_begin:self.base_end]
        if isinstance(piece, np.ndarray):
            piece = piece.tobytes()
        self.template = ctx.template_from_sequences([piece])
        self.ctx = ctx

    def scan(self, masters, mismatches, gather=True, group=None):
        """Hits of the whole template in whole-template coordinates (every rank
        gets all of them when gather=True)."""
        hits, stats = self.ctx.scan(self.template, masters, mismatches)
        lens = [len(m) for m in masters]
        shard_len = self.base_end - self.base_begin
        hits = localize_hits(hits, self.base_begin, shard_len, self.t_len, lens)
        hits = keep_owned(hits, self.start_begin, self.start_end, self.t_len, lens)
        if gather:
            hits = sort_hits(all_gather_records(hits, group))
        return hits, stats
