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


def shard_pairs(n_pairs, world_size, rank):
    """[begin, end) of the sequence pairs rank `rank` evaluates in a sharded `find_dimers_batch` /
    `SecStructures.batch`: pairs are independent, so the shards are contiguous, differ by
    at most one pair and need no exchange but the final gather of the (small) result records."""
    base, extra = divmod(n_pairs, world_size)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


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


def wire_to_hits(raw, primer_lens, rec_starts, rec_lens, rec_index=None):
    """Raw wire records (dp_wire_hit, include/degenprimer_b200.h) -> dp_hit records in
    the senders' record coordinates.  pattern = 2*primer + strand.  rec_starts /
    rec_lens are indexed by the record number of the wire record, or by rec_index
    (e.g. the sender rank of every record, when each rank holds one segment)."""
    out = np.zeros(len(raw), dtype=nat.HIT_DTYPE)
    if not len(raw):
        return out
    primer = (raw['pattern'] >> 1).astype(np.int64)
    strand = (raw['pattern'] & 1).astype(np.uint8)
    L = np.asarray(primer_lens, dtype=np.int64)[primer]
    rec = raw['record'].astype(np.int64) if rec_index is None else rec_index
    off = raw['g'].astype(np.int64) - np.asarray(rec_starts, dtype=np.int64)[rec]
    rl = np.asarray(rec_lens, dtype=np.int64)[rec]
    out['pos'] = np.where(strand == 0, off, rl - L - off).astype(np.uint64)
    out['record'] = raw['record']
    out['primer'] = primer.astype(np.uint16)
    out['strand'] = strand
    out['score3'] = raw['score3'].astype(np.uint8)
    return out


class HitExchange(object):
    """All-gather of the hits of every fused call, done by the call's own last kernel.

    Every rank owns two receive buffers that its peers map through CUDA IPC; the
    library's publish kernel stores {header, raw hits} of a call straight into the
    receive buffer of every rank over NVLink (no collective launch, no extra host
    round trip) and moves the previous call's slots to pinned host memory.
    torch.distributed only carries the IPC handles at set-up and the barrier at
    tear-down.  If any rank has more hits than a slot holds, `collect` returns None
    and the caller falls back to all_gather_records."""

    def __init__(self, ctx, max_hits=1024, group=None):
        import torch.distributed as dist
        self.ctx = ctx
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        mine = ctx.comm_export(self.world, self.rank, max_hits)
        blobs = [None] * self.world
        dist.all_gather_object(blobs, mine, group=group)
        ctx.comm_attach(blobs)
        dist.barrier(group=group)        # nobody stores into a buffer that is not mapped yet

    def close(self):
        import torch.distributed as dist
        self.ctx.synchronize()
        dist.barrier(group=self.group)   # peers may still be storing into this rank's buffers
        self.ctx.comm_destroy()

    def collect(self, age=0):
        """(raw wire records, sender rank of each) of a call, or None.  age 0 = the call
        that just returned, 1 = the one before it (already on the host, costs nothing)."""
        return self.ctx.gathered(age)

    def collect_hits(self, age=0):
        """Converted + sorted hits of every rank (record = sender rank), or None."""
        return self.ctx.gathered_hits(age)


def hit_keys(hits):
    """Sort key of dp_scan's order (primer, strand, record, pos) as one uint64."""
    return ((hits['primer'].astype(np.uint64) << np.uint64(57)) | (hits['strand'].astype(np.uint64) << np.uint64(56)) |
            (hits['record'].astype(np.uint64) << np.uint64(40)) | hits['pos'].astype(np.uint64))


def sort_hits(hits):
    """The order dp_scan uses: (primer, strand, record, pos)."""
    if len(hits) and int(hits['pos'].max()) < (1 << 40) and int(hits['record'].max()) < (1 << 16):
        return hits[np.argsort(hit_keys(hits), kind='stable')]
    return np.sort(hits, order=['primer', 'strand', 'record', 'pos'])


class ShardedTemplate(object):
    """This rank's slice of one long template, resident on this rank's GPU."""

    def __init__(self, ctx, text, rank, world_size, max_primer_len=nat.DP_MAX_PRIMER_LEN):
        """`text`: the whole template (str / bytes / uint8 array); only this
        rank's slice is uploaded."""
        self.t_len = len(text)
        self.rank, self.world_size = rank, world_size
        self.max_primer_len = max_primer_len
        self.base_begin, self.base_end, self.start_begin, self.start_end = \
            shard_bounds(self.t_len, world_size, rank, max_primer_len)
        piece = text[self.base_begin:self.base_end]
        if isinstance(piece, np.ndarray):
            piece = piece.tobytes()
        self.template = ctx.template_from_sequences([piece])
        self.ctx = ctx

    def find(self, query, cond, exchange=None, group=None):
        """Fused scan + duplex dG of this rank's slice (window starts restricted to the
        owned range on the device).  Returns (hits of the WHOLE template in
        whole-template coordinates, sorted; this rank's dimers; this rank's local
        hits the dimers index; stats, dG stats).  `exchange`: a HitExchange (NCCL
        gather queued on the same stream); without it hits are gathered afterwards."""
        lens = [len(m) for m in query.masters]
        lo = self.start_begin - self.base_begin
        hi = self.start_end - self.base_begin
        local, dimers, st, dst = self.ctx.find(self.template, query, cond=cond, seg_begin=lo, seg_end=hi)
        h = exchange.collect_hits() if exchange is not None else None
        if h is not None:
            sender = h['record'].astype(np.int64)
            begins, sizes = self._all_bounds()
            h['record'] = 0
            hits = sort_hits(localize_hits(h, begins[sender], sizes[sender], self.t_len, lens))
        else:
            mine = localize_hits(local, self.base_begin, self.base_end - self.base_begin, self.t_len, lens)
            hits = sort_hits(all_gather_records(mine, group))
        return hits, dimers, local, st, dst

    def _all_bounds(self):
        if getattr(self, '_bounds', None) is None:
            b = [shard_bounds(self.t_len, self.world_size, g, self.max_primer_len)[:2]
                 for g in range(self.world_size)]
            self._bounds = (np.array([x for x, _ in b], dtype=np.int64),
                            np.array([y - x for x, y in b], dtype=np.int64))
        return self._bounds

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


class GenomeLayout(object):
    """A genome of several records cut into `world` contiguous ranges of its padded
    concatenation (every record starts on a 32-base word), one range per rank — the
    strong-scaling split of SURVEY §8e.  Rank g owns the window STARTS in
    [ranges[g][0], ranges[g][1]); the bases it uploads are the pieces of the records that
    intersect that range, the last piece extended by a (max_primer_len - 1) halo so that every
    owned window is complete.  Only the last piece of a rank can carry a halo (the records
    behind it start past the range), so the owned starts of a rank are one interval
    [0, seg_end) of ITS padded concatenation — exactly what dp_find's seg_begin / seg_end take."""

    def __init__(self, lengths, world, max_primer_len):
        self.lengths = np.asarray(lengths, dtype=np.int64)
        words = (self.lengths + 31) // 32
        self.starts = (np.concatenate(([0], np.cumsum(words)[:-1])) * 32).astype(np.int64)
        self.n_words = int(words.sum())
        self.world = int(world)
        self.halo = int(max_primer_len) - 1
        per = -(-self.n_words // self.world) if self.n_words else 0
        self.ranges = [(min(self.n_words, g * per) * 32, min(self.n_words, (g + 1) * per) * 32)
                       for g in range(self.world)]
        self.pieces = [self._pieces(g) for g in range(self.world)]

    def _pieces(self, g):
        """[(record, offset, length, owned)]: bases [offset, offset+length) of the record are
        uploaded, window starts [offset, offset+owned) of it belong to rank g."""
        b, e = self.ranges[g]
        out = []
        for r in range(len(self.lengths)):
            s, n = int(self.starts[r]), int(self.lengths[r])
            lo, hi = max(0, b - s), min(n, e - s)
            if hi <= lo:
                continue
            out.append((r, lo, min(n, hi + self.halo) - lo, hi - lo))
        return out

    def records_of(self, g):
        return sorted({p[0] for p in self.pieces[g]})

    def local_starts(self, g):
        n = np.array([p[2] for p in self.pieces[g]], dtype=np.int64)
        return (np.concatenate(([0], np.cumsum((n + 31) // 32)[:-1])) * 32).astype(np.int64) if len(n) else n

    def seg_end(self, g):
        """Window starts of rank g's template that it owns: [0, seg_end)."""
        if not self.pieces[g]:
            return 0
        return int(self.local_starts(g)[-1]) + self.pieces[g][-1][3]

    def exchange_tables(self):
        """(base, local_start, offset, record, record_len) for Context.comm_set_layout."""
        base, ls, off, rec, rl = [0], [], [], [], []
        for g in range(self.world):
            starts = self.local_starts(g)
            for j, (r, o, _n, _own) in enumerate(self.pieces[g]):
                ls.append(int(starts[j]))
                off.append(o)
                rec.append(r)
                rl.append(int(self.lengths[r]))
            base.append(len(ls))
        return base, ls, off, rec, rl

    def to_global(self, hits, g, primer_lens):
        """Hits of rank g's template (record = piece index) in whole-genome coordinates."""
        out = hits.copy()
        if not len(hits):
            return out
        pc = self.pieces[g]
        rec = np.array([p[0] for p in pc], dtype=np.int64)[hits['record']]
        off = np.array([p[1] for p in pc], dtype=np.int64)[hits['record']]
        ln = np.array([p[2] for p in pc], dtype=np.int64)[hits['record']]
        L = np.asarray(primer_lens, dtype=np.int64)[hits['primer']]
        pos = hits['pos'].astype(np.int64)
        fwd = np.where(hits['strand'] == 0, pos, ln - L - pos) + off
        out['pos'] = np.where(hits['strand'] == 0, fwd, self.lengths[rec] - L - fwd).astype(np.uint64)
        out['record'] = rec.astype(np.uint32)
        return out


class ShardedGenome(object):
    """This rank's part of a multi-record genome, resident on its GPU (GenomeLayout)."""

    def __init__(self, ctx, layout, rank, record_source):
        """record_source(r) -> the bytes of whole record r: a pinned uint8 torch tensor / numpy
        array (sliced without a copy) or bytes."""
        self.ctx, self.layout, self.rank = ctx, layout, rank
        self._keep, srcs = [], []
        for r, off, n, _own in layout.pieces[rank]:
            src = record_source(r)
            if hasattr(src, 'data_ptr'):                      # torch tensor
                srcs.append((src.data_ptr() + off, n))
            elif isinstance(src, np.ndarray):
                srcs.append((src.ctypes.data + off, n))
            else:
                srcs.append(bytes(src[off:off + n]))
            self._keep.append(src)
        self.sources = srcs
        self.template = ctx.template_from_sequences(srcs)
        self.seg_end = layout.seg_end(rank)

    def attach(self, exchange):
        """Tells a HitExchange where the pieces of every rank come from: collect_hits() then
        returns the hits of the WHOLE genome in whole-genome coordinates."""
        exchange.ctx.comm_set_layout(*self.layout.exchange_tables())

    def find(self, query, cond, ctx=None, algo=0):
        """dp_find over the owned window starts.  Returns (local hits, dimers, stats, dG stats);
        GenomeLayout.to_global() translates the hits."""
        c = ctx or self.ctx
        return c.find(self.template, query, cond=cond, seg_begin=0, seg_end=self.seg_end, algo=algo)
