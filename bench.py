#!/usr/bin/env python3
"""bench.py — the hot path's headline metric on B200 (contract: task brief ④).

metric : template bp x primer-variants scanned per second (BASELINE.json), with the
         dG stage run on every hit inside the same step.
step   : one pass of the hot path over the resident genome: the scan of every
         primer over both strands + the duplex dG of every (hit, variant), one
         stream submission (dp_find), results back on the host.
workload (default): BASELINE.json configs[3], the largest single-GPU config —
         `config4_20plex_3Gbp`: 20-primer multiplex (106 variants) vs a synthetic
         3.1 Gbp, 24-record genome, k = 3.  The north-star pair (256-fold degenerate,
         k = 3) is timed on the SAME resident genome and reported as `north_star`.
         `--workload config2_tp53_chr17` etc. select the other configs.
N > 1  : STRONG scaling — the one genome is cut into N contiguous ranges of its
         padded concatenation with an (L_max-1) halo (distributed.GenomeLayout),
         rank g scans range g, and every rank ends each step with the hit set of
         the whole genome (peer-memory stores from the call's last kernel).
         After the timed region the gathered set is compared with the hits of
         the whole genome scanned by rank 0 alone.
CPU    : `cpu_baseline` (N = 1) and `--impl reference` time the reference's
         algorithm (oracle O1: scipy FFT scan + Duplex objects) on all host cores
         over the SAME bounded sample of the workload every time
         (synthetic.cpu_sample_segments); the GPU's hits and dimers on those very
         bytes are compared with it (`parity`), and three seeded 50 Mbp stretches
         are re-scanned by the C brute-force oracle.

`python bench.py --gpus N --steps K --warmup W [--impl reference]`
prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'template_bp_x_primer_variants_scanned_per_s'
UNIT = 'bp*variants/s'
WORKLOAD = 'config4_20plex_3Gbp'
NORTH_STAR = 'northstar_256fold_3Gbp'
COND = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0)     # reference README.md:72-96
L2_BYTES = 126 * 1024 * 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default=WORKLOAD)
    ap.add_argument('--scale', type=float, default=1.0, help='shrink the template (testing only)')
    ap.add_argument('--algo', type=int, default=0, help='0 auto, 1 popcount, 2 bit-slice, 3 run-time specialised')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-parity', action='store_true', help='skip the C brute-force re-scan of the 50 Mbp stretches')
    ap.add_argument('--no-north-star', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-jit', action='store_true', help='generic scan kernel instead of the run-time specialised one')
    ap.add_argument('--parity-mbp', type=float, default=50.0)
    ap.add_argument('--cpu-repeats', type=int, default=2)
    ap.add_argument('--host-threads', type=int, default=3, help='host threads that queue and collect the calls')
    ap.add_argument('--in-flight', type=int, default=6,
                    help='calls in flight on the GPU (one context per call); 1 = strictly sequential')
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    return rank, world, local


def variants_of(primers):
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    return [unambiguous_sequences(p) for p in primers]


class Plan(object):
    """What both arms need to know about a workload; `config` is the same dict in both."""

    def __init__(self, name, scale):
        from degenprimer_b200 import synthetic
        spec = synthetic._spec(name)
        self.name, self.scale = name, scale
        self.primers, self.k = list(spec['primers']), spec['k']
        self.var_lists = variants_of(self.primers)
        self.n_var = sum(len(v) for v in self.var_lists)
        self.lengths = synthetic.record_lengths(name, scale)
        self.bp = int(sum(self.lengths))
        self.max_len = max(len(p) for p in self.primers)
        self.config = {
            'workload': name, 'template_bp': self.bp, 'records': len(self.lengths),
            'primers': len(self.primers), 'primer_len': [min(map(len, self.primers)), self.max_len],
            'variants': self.n_var, 'max_mismatches': self.k, 'scale': scale,
            'cpu_sample': 'synthetic.cpu_sample_segments: %d disjoint stretches of record 0 around planted sites, '
                          '%d bp x primers per pass' % (synthetic.CPU_SAMPLE_SEGMENTS, synthetic.CPU_SAMPLE_BP_PRIMERS),
            'inputs': 'synthetic genome, 0.375 B/bp resident (%.0f MB of planes > 126 MB L2)' % (0.375 * self.bp / 1e6),
        }


# ---- clocks -------------------------------------------------------------------------
class ClockSampler(object):
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md) through
    NVML in a background thread (every 5 ms; the nvidia-smi CLI takes longer to start than the
    timed region lasts)."""
    REASONS = (('hw_slowdown', 0x8), ('hw_thermal_slowdown', 0x40), ('sw_thermal_slowdown', 0x20),
               ('sw_power_cap', 0x4))

    def __init__(self, device):
        self.device = device
        self.thread = None
        self.samples, self.maxes, self.bits = [], [], 0
        self._stop = False

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get('CUDA_VISIBLE_DEVICES')
            index = int(visible.split(',')[self.device]) if visible else self.device
            self.nv = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
        except Exception:
            self.nv = None
            return

        def loop():
            nv, h = self.nv, self.handle
            while not self._stop:
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                    self.maxes.append(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
                    self.bits |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    try:
                        self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                    except Exception:
                        pass
                time.sleep(0.005)
        self.thread = threading.Thread(target=loop, daemon=True)
        self.thread.start()

    def stop(self):
        out = {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
        if self.thread is None:
            return out
        self._stop = True
        self.thread.join(timeout=2)
        if self.samples:
            out.update(sm_mhz=float(np.median(self.samples)), sm_max_mhz=float(max(self.maxes)) if self.maxes else None,
                       samples=len(self.samples))
        out['reasons'] = [name for name, bit in self.REASONS if self.bits & bit]
        return out


class HostThreads(object):
    """A few long-lived host threads, each running the callables handed to it one after the other."""

    def __init__(self, n):
        import queue
        self.todo = [queue.Queue() for _ in range(n)]
        self.done = queue.Queue()
        self.threads = [threading.Thread(target=self._loop, args=(i,), daemon=True) for i in range(n)]
        for th in self.threads:
            th.start()

    def _loop(self, i):
        while True:
            fn, a = self.todo[i].get()
            if fn is None:
                return
            try:
                fn(*a)
            finally:
                self.done.put(i)

    def run(self, i, fn, *a):
        self.todo[i].put((fn, a))

    def wait(self, n):
        for _ in range(n):
            self.done.get()

    def close(self):
        for q in self.todo:
            q.put((None, ()))


# ---- CPU arm: the reference algorithm (oracle O1) on the host cores -------------------------
# Only the cpu_baseline / parity legs and --impl reference call into oracle/.
_CPU_STATE = {}


def _mask_of(positions):
    m = 0
    for i in positions:
        m |= 1 << int(i)
    return m


def _cpu_task(job):
    """One (stretch, primer) of the sample in a fork worker: SearchEngine.find of the reference =
    FFT scan of both strands (its slices run in this process) + a Duplex per (site, variant)."""
    from oracle import conditions as ocond
    from oracle import search_o1 as o1
    from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord
    seg, pi = job
    st = _CPU_STATE
    text, master = st['texts'][seg], st['primers'][pi]
    m = o1.find_matches(text, master, st['k'], workers=0)
    fwd_seq, rev_seq, t_len, p_len, fm, rm = m
    primer = OPrimer(OSeqRecord(master, id='p%d' % pi), 1.5e-6)
    sites = []
    with ocond.AcquireParameters():
        for strand, seq, matches in ((0, fwd_seq, fm), (1, rev_seq, rm)):
            for pos in matches:
                _, dups = o1.compile_duplexes_for_position(int(pos), seq, primer, t_len, p_len, bool(strand))
                sites.append((strand, int(pos),
                              [(d._fwd_sequence, [(_mask_of(x.fwd_matches), float(x.dG)) for x in d.dimers])
                               for d in dups]))
    return seg, pi, sites


class CpuArm(object):
    """The bounded sample of a workload and a fork pool (made BEFORE any CUDA / OpenMP use) that
    scans it with the reference algorithm."""

    def __init__(self, plan):
        from degenprimer_b200 import synthetic
        from oracle import conditions as ocond
        import multiprocessing
        ocond.PCR_P.set(COND)
        self.plan = plan
        self.segments, self.record0 = synthetic.cpu_sample_segments(plan.name, plan.scale)
        self.sample_bp = sum(b - a for _, a, b in self.segments)
        _CPU_STATE.update(texts=[self.record0[a:b].tobytes().decode() for _, a, b in self.segments],
                          primers=plan.primers, k=plan.k)
        self.cores = os.cpu_count() or 1
        self.jobs = [(s, p) for s in range(len(self.segments)) for p in range(len(plan.primers))]
        self.pool = multiprocessing.get_context('fork').Pool(self.cores) if self.cores > 1 else None

    def run(self):
        """One pass over the sample.  Returns (seconds, results)."""
        t0 = time.perf_counter()
        if self.pool is not None:
            res = self.pool.map(_cpu_task, self.jobs, chunksize=1)
        else:
            res = [_cpu_task(j) for j in self.jobs]
        return time.perf_counter() - t0, res

    def close(self):
        if self.pool is not None:
            self.pool.close()
            self.pool.join()
            self.pool = None
        _CPU_STATE.clear()

    def describe(self, seconds, res):
        hits = sum(len(s) for _, _, s in res)
        dimers = sum(len(dm) for _, _, s in res for _, _, dups in s for _, dm in dups)
        return dict(value=self.sample_bp * self.plan.n_var / seconds, unit=UNIT, cores=self.cores, kind='port',
                    sample='%d stretches of record 0 of %s = %d bp x %d primers (%d variants): FFT scan of both '
                           'strands (scipy.fftpack, %d fork workers over stretch x primer) + a Duplex per '
                           '(site, variant): %d sites, %d dimers; %.2f s per pass'
                           % (len(self.segments), self.plan.name, self.sample_bp, len(self.plan.primers),
                              self.plan.n_var, self.cores, hits, dimers, seconds),
                    seconds=seconds, hits=hits, dimers=dimers, sample_bp=self.sample_bp,
                    extrapolation='whole workload = %.1f x the sample (throughput is per bp)'
                                  % (self.plan.bp / max(1, self.sample_bp)))


def run_reference(args):
    """--impl reference: the reference's own CPU algorithm for this path.  The reference itself
    is Python 2.7 + Biopython and cannot run in this image (no python2, no Bio), so
    its Python-3 restatement (oracle O1) is what is timed; `kind` says so.  Every step scans the
    same bounded sample of the workload the GPU arm's cpu_baseline leg scans."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    plan = Plan(args.workload, args.scale)
    arm = CpuArm(plan)
    for _ in range(min(args.warmup, 1)):          # imports, page faults; more warm-up adds nothing on a CPU
        arm.run()
    t0 = time.perf_counter()
    res = None
    for _ in range(max(1, args.steps)):
        _dt, res = arm.run()
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    arm.close()
    cpu = arm.describe(dt, res)
    value = cpu['value']
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True,
        'scaling': 'strong', 'vs_baseline': None, 'dtype': 'complex128 FFT + f64 dG (reference algorithm)',
        'data': 'synthetic', 'config': plan.config, 'cpu_baseline': cpu,
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line))


def bind_near_gpu(local):
    """Pins this rank to the host cores NVML lists as local to its GPU, so that pinned staging
    memory is first touched on the GPU's own NUMA node."""
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get('CUDA_VISIBLE_DEVICES')
        index = int(visible.split(',')[local]) if visible else local
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        n_words = (os.cpu_count() + 63) // 64
        masks = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = {64 * w + b for w, m in enumerate(masks) for b in range(64) if (int(m) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return 0


# ---- parity helpers ------------------------------------------------------------------------------
def gpu_sites_in(hits, dimers, record, a, b, primer_lens, rec_len):
    """{(primer, strand, pos in the stretch's own coordinates): [(variant, mask, dG), ...]} of the GPU
    hits whose window lies inside bases [a, b) of `record` (strand 1: position on the reverse
    complement of the stretch, as the CPU arm sees it)."""
    L = np.asarray(primer_lens, dtype=np.int64)[hits['primer']]
    pos = hits['pos'].astype(np.int64)
    fwd = np.where(hits['strand'] == 0, pos, rec_len - L - pos)
    sel = np.flatnonzero((hits['record'] == record) & (fwd >= a) & (fwd + L <= b))
    bounds = np.searchsorted(dimers['hit'], np.arange(len(hits) + 1)) if dimers is not None else None
    out = {}
    for h in sel:
        local = int(fwd[h] - a)
        p = local if hits['strand'][h] == 0 else (b - a) - int(L[h]) - local
        rows = dimers[bounds[h]:bounds[h + 1]] if dimers is not None else ()
        out[(int(hits['primer'][h]), int(hits['strand'][h]), p)] = \
            [(int(r['variant']), int(r['match_mask']), float(r['dG'])) for r in rows]
    return out


def compare_with_cpu(arm, res, hits, dimers, plan, rec_len):
    """GPU results on the sample's bytes against the CPU arm's.  Returns the parity record."""
    want = {}
    for seg, pi, sites in res:
        for strand, pos, dups in sites:
            rows = []
            for vseq, dm in dups:
                v = plan.var_lists[pi].index(vseq)
                rows.extend((v, m, g) for m, g in dm)
            want[(seg, pi, strand, pos)] = rows
    got = {}
    lens = [len(p) for p in plan.primers]
    for s, (rec, a, b) in enumerate(arm.segments):
        for (pi, strand, pos), rows in gpu_sites_in(hits, dimers, rec, a, b, lens, rec_len).items():
            got[(s, pi, strand, pos)] = rows
    equal = set(got) == set(want)
    n_dim, worst = 0, 0.0
    for key in set(got) & set(want):
        g, w = got[key], want[key]
        if [(v, m) for v, m, _ in g] != [(v, m) for v, m, _ in w]:
            equal = False
            continue
        for (_, _, x), (_, _, y) in zip(g, w):
            worst = max(worst, abs(x - y))
            n_dim += 1
    if worst > 0.01:
        equal = False
    return {'cpu_sample_hits_gpu': len(got), 'cpu_sample_hits_cpu': len(want), 'dimers_checked': n_dim,
            'dG_max_abs_diff': worst, 'dG_tolerance': 0.01, 'equal': bool(equal)}


def all_equal(d):
    """Every 'equal' of a nested parity record."""
    ok = bool(d.get('equal', True))
    for v in d.values():
        if isinstance(v, dict):
            ok = ok and all_equal(v)
    return ok


def brute_force_check(plan, records, segments, hits):
    """The C brute-force oracle over seeded stretches of the genome, every primer, both strands,
    against the GPU's hits whose window lies inside the stretch.  `records[r]` = uint8 array."""
    from oracle import o2c
    lens = [len(p) for p in plan.primers]
    checked, ok = 0, True
    t0 = time.perf_counter()
    for rec, a, b in segments:
        text = records[rec][a:b].tobytes()
        got = gpu_sites_in(hits, None, rec, a, b, lens, len(records[rec]))
        want = set()
        for pi, master in enumerate(plan.primers):
            f, r = o2c.find(text, master, plan.k)
            want.update((pi, 0, int(x)) for x in f)
            want.update((pi, 1, int(x)) for x in r)
        checked += len(want)
        if set(got) != want:
            ok = False
    return {'segments': [[int(r), int(a), int(b)] for r, a, b in segments], 'hits_checked': checked,
            'equal': bool(ok), 'seconds': time.perf_counter() - t0,
            'oracle': 'oracle/o2_scan.c (brute-force thirds score of every window, both strands)'}


# ---- GPU arm -----------------------------------------------------------------------------------
def run_ours(args):
    rank, world, local = dist_env()
    if world > 1:
        bind_near_gpu(local)
    if world != args.gpus and world > 1:
        raise SystemExit('WORLD_SIZE=%d but --gpus %d' % (world, args.gpus))
    from degenprimer_b200 import synthetic
    from degenprimer_b200.distributed import GenomeLayout, ShardedGenome, sort_hits
    plan = Plan(args.workload, args.scale)
    with_ns = (args.workload == WORKLOAD and not args.no_north_star)
    ns_plan = Plan(NORTH_STAR, args.scale) if with_ns else None
    if ns_plan is not None and ns_plan.lengths != plan.lengths:
        raise SystemExit('the north-star pair and %s do not share a genome' % args.workload)
    max_len = max(plan.max_len, ns_plan.max_len if ns_plan else 0)
    layout = GenomeLayout(plan.lengths, world, max_len)

    # ---- CPU legs first: a fork pool must exist before CUDA / OpenMP threads do -------------
    cpu = cpu_res = arm = None
    ns_cpu = ns_res = ns_arm = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        arm = CpuArm(plan)
        arm.run()                                   # imports, page faults
        best = None
        for _ in range(max(1, args.cpu_repeats)):
            dt, cpu_res = arm.run()
            best = dt if best is None else min(best, dt)
        arm.close()
        cpu = arm.describe(best, cpu_res)
        if ns_plan is not None:
            ns_arm = CpuArm(ns_plan)
            dt, ns_res = ns_arm.run()
            ns_arm.close()
            ns_cpu = ns_arm.describe(dt, ns_res)

    # ---- the genome: every rank generates the records its range touches ------------------------
    t0 = time.perf_counter()
    need = set(layout.records_of(rank))
    par_segs = []
    if rank == 0 and not args.no_parity:
        par_segs = synthetic.parity_segments(plan.name, plan.scale, seg_bp=int(args.parity_mbp * 1e6))
    whole_check = rank == 0 and world > 1            # rank 0 also holds the whole genome for the N > 1 check
    if whole_check:
        need = set(range(len(plan.lengths)))
    if world == 1:
        need |= {r for r, _, _ in par_segs}
    wl = synthetic.make_workload(plan.name, scale=plan.scale, only_records=sorted(need))
    s_generate = time.perf_counter() - t0

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    os.environ['DEGENPRIMER_B200_DEVICE'] = str(local)
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import _native as nat
    tdf.PCR_P.set(COND)
    ctx = tdf._ctx()
    info = ctx.device_info()
    cond = tdf.conditions()

    # host records page-locked in place (no second copy of 3 GB): H2D runs at PCIe speed
    cudart = torch.cuda.cudart()
    pinned = []
    for r in sorted(need):
        arr = wl['records'][r]
        if arr is not None and len(arr):
            try:
                if int(cudart.cudaHostRegister(arr.ctypes.data, arr.nbytes, 0)) == 0:
                    pinned.append(arr)
            except Exception:
                pass
    t0 = time.perf_counter()
    shard = ShardedGenome(ctx, layout, rank, lambda r: wl['records'][r])
    s_upload = time.perf_counter() - t0
    tmpl = shard.template
    plane_bytes = tmpl.device_bytes
    shard_bp = int(sum(p[3] for p in layout.pieces[rank]))

    # Lanes: contexts (own stream, own result buffers) on this GPU, one call in flight per lane, so
    # that the host work of a call (queueing, result conversion) overlaps the kernels of another
    n_lanes = max(1, args.in_flight)
    lanes = [ctx] + [tdf.new_context(local) for _ in range(n_lanes - 1)]
    exchs = [None] * n_lanes
    if world > 1:
        # the hit gather is the epilogue of the publish kernel: each rank stores its slot into
        # every peer's receive buffer over NVLink (CUDA IPC peer memory), no collective launch
        from degenprimer_b200.distributed import HitExchange
        exchs = [HitExchange(c, max_hits=1 << 16) for c in lanes]
        for ex in exchs:
            shard.attach(ex)

    def sync_all():
        for c in lanes:
            c.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # CUDA events on the stream the kernels are launched on (lane 0's stream; when the region
    # ends every lane has been waited for, so the closing event is later than all work)
    kstream = torch.cuda.ExternalStream(ctx.stream, device=torch.device('cuda', local))

    def drive(queries, my, n_steps, acc):
        """n_steps steps on the lanes `my` (indices), one call in flight per lane, from one host
        thread.  At N > 1 a step ends with the hit set of the WHOLE genome on this rank: the
        call two back is collected (converted, sorted) while the newest one runs."""
        F = len(my)
        made = [0] * F

        clock = time.perf_counter

        def begin(i):
            j = i % F
            t = clock()
            lanes[my[j]].find_begin(tmpl, queries[my[j]], cond, seg_begin=0, seg_end=shard.seg_end, algo=args.algo)
            acc['host_begin'] += clock() - t
            made[j] += 1

        for i in range(min(F - 1, n_steps)):
            begin(i)
        for i in range(n_steps):
            if i + F - 1 < n_steps:
                begin(i + F - 1)
            j = i % F
            lane = my[j]
            t = clock()
            if exchs[lane] is not None and made[j] >= 3:
                acc['hits'] += len(exchs[lane].collect_hits(2))
            t1 = clock()
            hits, dimers, st, dst = lanes[lane].find_end()
            acc['host_collect'] += t1 - t
            acc['host_end'] += clock() - t1
            if exchs[lane] is None:
                acc['hits'] += len(hits)
            acc['own_hits'] += len(hits)
            acc['ms_main'] += st['ms_main']
            acc['ms_scan'] += st['ms_scan']
            acc['ms_dg'] += dst['ms_kernel']
            acc['ms_total'] += st['ms_total']
            acc['launches'] += st['launches'] + dst['launches'] + 1        # + the publish kernel
            acc['dimers'] += len(dimers)
            acc['duplexes'] += dst['duplexes']
            acc['cand'] += st['candidates']
            acc['windows'] = st['windows']
            acc['algo'] = st['algo']
        for j in range(F):
            lane = my[j]
            if exchs[lane] is not None:
                if made[j] >= 2:
                    acc['hits'] += len(exchs[lane].collect_hits(1))
                if made[j] >= 1:
                    acc['hits'] += len(exchs[lane].collect_hits(0))

    helpers = HostThreads(max(0, min(args.host_threads, n_lanes) - 1))

    def timed_region(queries, K, F, T):
        """K steps with F calls in flight (one per lane), queued and collected by T host threads
        (F = T = 1: strictly one call after the other on the main thread)."""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        T = max(1, min(T, F))
        accs = [dict(ms_main=0.0, ms_scan=0.0, ms_dg=0.0, ms_total=0.0, launches=0, hits=0, own_hits=0, dimers=0, duplexes=0,
                     cand=0, windows=0, algo=0, host_begin=0.0, host_collect=0.0, host_end=0.0) for _ in range(T)]
        shares = [K // T + (1 if t < K % T else 0) for t in range(T)]
        groups = [list(range(t, F, T)) for t in range(T)]
        for c in lanes:
            c.set_timing(F == 1)               # the ms_* stats are only read in the sequential region
        errors = []

        def work(t):
            try:
                drive(queries, groups[t], shares[t], accs[t])
            except BaseException as e:          # noqa: surfaced on the main thread below
                errors.append(e)

        sync_all()
        t0 = time.perf_counter()
        ev0.record(kstream)
        # the host threads live as long as the bench (a caller's thread pool): a NEW thread pays the
        # CUDA runtime's per-thread set-up on its first call, ~10 ms next to 7 other ranks doing the same
        for t in range(1, T):
            helpers.run(t - 1, work, t)
        work(0)
        helpers.wait(T - 1)
        if errors:
            raise errors[0]
        for c in lanes:
            c.synchronize()
        ev1.record(kstream)
        sync_all()
        wall = time.perf_counter() - t0
        elapsed = ev0.elapsed_time(ev1) * 1e-3        # device clock, first launch -> last result on the host
        if world > 1:
            tt = torch.tensor([elapsed, wall], dtype=torch.float64, device='cuda')
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            elapsed, wall = float(tt[0].item()), float(tt[1].item())
        acc = accs[0]
        for o in accs[1:]:
            for key in ('ms_main', 'ms_scan', 'ms_dg', 'ms_total', 'launches', 'hits', 'own_hits', 'dimers', 'duplexes', 'cand',
                        'host_begin', 'host_collect', 'host_end'):
                acc[key] += o[key]
        return elapsed, wall, acc

    def measure(p, K, W):
        """Warm-up, the pipelined headline region, then the same steps strictly sequentially (per-call
        latency and kernel durations that no other lane's kernels stretch)."""
        queries = [c.prepare(p.primers, p.k, p.var_lists, jit=not args.no_jit) for c in lanes]
        n_threads = max(1, min(args.host_threads, n_lanes))
        timed_region(queries, max(W, 3 * n_lanes), n_lanes, n_threads)     # every lane: workspaces, page-locked result blocks
        elapsed, wall, acc = timed_region(queries, K, n_lanes, n_threads)
        ks = max(3, min(K, 10))
        seq_elapsed, _w, seq = timed_region(queries, ks, 1, 1)
        return dict(queries=queries, elapsed=elapsed, wall=wall, acc=acc, K=K, seq_elapsed=seq_elapsed, seq=seq, ks=ks)

    K = max(1, args.steps)
    W = max(3, args.warmup)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    main = measure(plan, K, W)
    ns = measure(ns_plan, K, W) if ns_plan is not None else None
    clocks = sampler.stop() if rank == 0 else None

    # ---- parity, outside the timed regions ----------------------------------------------------
    parity = None
    lens = [len(p) for p in plan.primers]
    if world == 1 and rank == 0:
        hits, dimers, _st, _dst = shard.find(main['queries'][0], cond, algo=args.algo)
        hits_g = layout.to_global(hits, 0, lens)           # N = 1: pieces are the records themselves
        parity = {}
        if arm is not None:
            parity.update(compare_with_cpu(arm, cpu_res, hits_g, dimers, plan, len(arm.record0)))
        if par_segs:
            parity['brute_force'] = brute_force_check(plan, wl['records'], par_segs, hits_g)
        if ns_plan is not None and ns_arm is not None:
            h2, d2, _s2, _d2 = shard.find(ns['queries'][0], cond, algo=args.algo)
            ns_par = compare_with_cpu(ns_arm, ns_res, layout.to_global(h2, 0, [len(p) for p in ns_plan.primers]),
                                      d2, ns_plan, len(ns_arm.record0))
            if par_segs:
                ns_par['brute_force'] = brute_force_check(ns_plan, wl['records'], par_segs[:1],
                                                          layout.to_global(h2, 0, [len(p) for p in ns_plan.primers]))
            parity['north_star'] = ns_par
        parity['equal_all'] = all_equal(parity)
    elif world > 1:
        # one more step: every rank's gathered, converted hit set of the whole genome ...
        lanes[0].find_begin(tmpl, main['queries'][0], cond, seg_begin=0, seg_end=shard.seg_end, algo=args.algo)
        lanes[0].find_end()
        gathered = exchs[0].collect_hits(0)
        same = torch.tensor([1], dtype=torch.int64, device='cuda')
        ref_n = torch.tensor([0], dtype=torch.int64, device='cuda')
        if rank == 0:
            # ... against the whole genome scanned by rank 0 alone
            one = GenomeLayout(plan.lengths, 1, max_len)
            whole = ShardedGenome(lanes[0], one, 0, lambda r: wl['records'][r])
            h, _d, _s, _ds = lanes[0].find(whole.template, main['queries'][0], cond=cond, algo=args.algo)
            ref = sort_hits(one.to_global(h, 0, lens))
            whole.template.close()
            ok = len(ref) == len(gathered) and all(
                np.array_equal(ref[f], gathered[f]) for f in ('pos', 'record', 'primer', 'strand', 'score3'))
            same[0] = 1 if ok else 0
            ref_n[0] = len(ref)
        # every rank must hold the same bytes as rank 0's gathered set
        digest = int(np.frombuffer(gathered.tobytes(), dtype=np.uint8).astype(np.uint64).sum()) if len(gathered) else 0
        dg = torch.tensor([digest, len(gathered)], dtype=torch.int64, device='cuda')
        dgs = [torch.zeros_like(dg) for _ in range(world)]
        dist.all_gather(dgs, dg)
        dist.broadcast(same, 0)
        dist.broadcast(ref_n, 0)
        parity = {'gathered_hits': int(len(gathered)), 'whole_genome_hits_rank0': int(ref_n.item()),
                  'gathered_equals_whole_genome': bool(same.item()),
                  'same_on_every_rank': bool(all(torch.equal(x, dgs[0]) for x in dgs)),
                  'equal_all': bool(same.item()) and bool(all(torch.equal(x, dgs[0]) for x in dgs))}
    for ex in exchs:
        if ex is not None:
            ex.close()          # the e2e leg below runs the ranks independently

    # ---- e2e: packed genome in pinned HOST memory -> reference-facing call -> host results ------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, plan, shard, lanes, sync_all, rank, world)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ----------------------------------------------------
    peaks = {}
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as fh:
            peaks = json.load(fh)
    except Exception:
        pass
    hbm_peak = float(peaks.get('hbm_gbs', 6650.0))
    peak_src = 'MEASURED_PEAKS.json hbm_gbs (burst copy)' if peaks else 'fallback 6.65 TB/s (B200_PROFILING.md)'
    lop3 = ctx.calibrate('lop3')
    popc = ctx.calibrate('popc')
    lds = ctx.calibrate('lds')
    prof = {}
    try:
        with open(os.path.join(ROOT, 'profiles', 'kernel_counters.json')) as fh:
            prof = json.load(fh)
    except Exception:
        pass

    def rooflines(p, m):
        seq, ks = m['seq'], m['ks']
        ms_main = seq['ms_main'] / ks
        algo_bytes = 0.25 * shard_bp                 # 2 bits per template base, read once
        achieved = algo_bytes / (ms_main * 1e-3) / 1e9 if ms_main > 0 else 0.0
        kernel = {2: 'dp_scan_bitslice_kernel', 3: 'dp_jit_scan'}.get(seq['algo'], 'dp_scan_popc_kernel')
        entry = prof.get(p.name, {}).get(kernel) if args.scale == 1.0 and world == 1 else None
        roof = {'bound': 'hbm', 'achieved': achieved, 'peak': hbm_peak, 'unit': 'GB/s',
                'frac': achieved / hbm_peak, 'traffic': entry.get('dram_bytes_per_launch') if entry else None,
                'traffic_source': entry.get('source') if entry else None,
                'kernel': kernel, 'ms_per_launch': ms_main, 'algorithmic_bytes_per_launch': algo_bytes,
                'launches_per_step': None, 'peak_source': peak_src,
                'note': 'integer-pipe bound by design (XOR/AND/LOP3 over %d patterns per word): see int_roofline'
                        % (2 * len(p.primers))}
        windows = float(seq['windows'])
        a_win = 16.0
        introof = {
            'bound': 'int32 ALU (LOP3) issue rate, measured by dp_calibrate on this GPU',
            'peak_lop3_lane_ops_per_s': lop3, 'peak_popc_lane_ops_per_s': popc, 'peak_lds32_lane_ops_per_s': lds,
            'windows_per_s': windows / (ms_main * 1e-3) if ms_main > 0 else 0.0,
            'canonical_ops_per_window': a_win,
            'frac_of_canonical_int_roof': (windows * a_win / (ms_main * 1e-3)) / lop3 if ms_main > 0 else 0.0,
            'note': 'canonical = 16 int32 ops incl. 2 POPC per window (SURVEY 8d); the bit-sliced prefilter issues '
                    'far fewer, so a fraction above 1 means "faster than the per-window popcount roof"',
        }
        if entry and entry.get('warp_inst_per_bp') and ms_main > 0:
            # instructions actually issued (ncu smsp__inst_executed.sum of the committed capture, per
            # template base) x 32 lanes / this run's kernel time, against the LOP3 rate calibrated in this run;
            # every instruction counts, so a kernel that sits on the ALU pipe reads close to (or a little
            # above) 1: its few non-ALU instructions issue on other pipes
            introof['alu_pipe_pct_ncu'] = entry.get('alu_pipe_pct')
            introof['frac_issued_source'] = entry.get('source')
            captured_ms = float(entry.get('ms_under_ncu') or 0.0)
            if captured_ms and ms_main < 0.95 * captured_ms:
                # this run's kernel is clearly shorter than the captured one: the kernel has changed since the
                # capture (fewer instructions), so the capture's instruction count no longer describes it
                introof['counters_stale'] = ('kernel time %.3f ms against %.3f ms in the committed capture: the '
                                             'instruction count of that capture is not this kernel\'s' % (ms_main, captured_ms))
            else:
                introof['warp_inst_per_launch'] = entry['warp_inst_per_bp'] * shard_bp
                introof['frac_issued'] = entry['warp_inst_per_bp'] * shard_bp * 32.0 / (ms_main * 1e-3) / lop3
                if entry.get('alu_pipe_pct') and captured_ms:
                    # the pipe this kernel is bound by: utilisation in the committed capture, scaled by that
                    # capture's kernel time over this run's (same kernel, same input)
                    introof['frac_alu_pipe'] = entry['alu_pipe_pct'] / 100.0 * captured_ms / ms_main
        return roof, introof

    def block(p, m):
        acc, seq, Kb, ks = m['acc'], m['seq'], m['K'], m['ks']
        ms_step = m['elapsed'] / Kb * 1e3
        return {
            'value': p.bp * p.n_var / (m['elapsed'] / Kb), 'ms_per_step': ms_step, 'steps': Kb,
            'wall_ms_per_step': m['wall'] / Kb * 1e3,
            'hits_per_step': acc['hits'] / Kb, 'duplexes_per_step': world * acc['duplexes'] / Kb if world == 1 else None,
            'dimers_per_step_rank0': acc['dimers'] / Kb, 'candidates_per_step_rank0': acc['cand'] / Kb,
            'ms_scan_kernels': seq['ms_scan'] / ks, 'ms_dg_kernels': seq['ms_dg'] / ks,
            'ms_call_on_device': seq['ms_total'] / ks,
            'sequential_ms_per_step': m['seq_elapsed'] / ks * 1e3,
            'dg_duplexes_per_s': (seq['duplexes'] / ks) / (seq['ms_dg'] / ks * 1e-3) if seq['ms_dg'] > 0 else None,
            'gpu_launches': acc['launches'],
            # where the host threads of rank 0 spent the pipelined region (summed over the threads, per step)
            'host_ms_per_step': {'find_begin': acc['host_begin'] / Kb * 1e3, 'collect_gathered': acc['host_collect'] / Kb * 1e3,
                                 'find_end_incl_wait': acc['host_end'] / Kb * 1e3},
        }

    mb = block(plan, main)
    roofline, int_roofline = rooflines(plan, main)
    line = {
        'metric': METRIC, 'value': mb['value'], 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': mb['ms_per_step'], 'higher_is_better': True, 'scaling': 'strong',
        'vs_baseline': None, 'dtype': 'u32 bitwise (scan) + f64 (dG)', 'data': 'synthetic',
        'config': plan.config,
        'run': {'calls_in_flight': n_lanes, 'host_threads': max(1, min(args.host_threads, n_lanes)),
                'timing': 'CUDA events on the context stream around K steps, %d calls in flight (one context each); max over '
                          'ranks; wall clock %.3f ms/step' % (n_lanes, mb['wall_ms_per_step']),
                'l2': 'input planes per rank %.0f MB (> 126 MB L2): every step re-reads them from HBM' % (plane_bytes / 1e6)
                      if plane_bytes > 1.5 * L2_BYTES else
                      'input planes per rank %.0f MB against %.0f MB of L2 (the scan streams the two base planes, %.0f MB, per '
                      'step and %d calls in flight read them side by side); no flush between steps: the kernel is bound by '
                      'the integer ALU, not by memory (DRAM 2 %% of peak in the ncu capture)'
                      % (plane_bytes / 1e6, L2_BYTES / 1e6, plane_bytes * 2 / 3e6, n_lanes),
                'sharding': 'one genome, %d contiguous ranges of the padded concatenation, halo %d bases, hit set of the '
                            'whole genome on every rank after every step' % (world, max_len - 1) if world > 1 else 'single GPU',
                'gpu': info['name'], 'sms': info['sm_count'], 's_generate': s_generate, 's_upload_pack': s_upload,
                'host_ms_per_step': mb['host_ms_per_step'], 'host_cores': os.cpu_count()},
        'hits_per_step': mb['hits_per_step'], 'dimers_per_step': mb['dimers_per_step_rank0'],
        'duplexes_per_step': mb['duplexes_per_step'], 'candidates_per_step': mb['candidates_per_step_rank0'],
        'ms_scan_kernels': mb['ms_scan_kernels'], 'ms_dg_kernels': mb['ms_dg_kernels'],
        'dg_duplexes_per_s': mb['dg_duplexes_per_s'], 'gpu_launches': mb['gpu_launches'],
        'sequential': {'ms_per_step': mb['sequential_ms_per_step'], 'ms_call_on_device': mb['ms_call_on_device'],
                       'note': 'the same step with one call in flight (per-call latency); the kernel durations of the '
                               'roofline are taken here'},
        'roofline': roofline, 'int_roofline': int_roofline, 'clocks': clocks, 'parity': parity,
    }
    if e2e is not None:
        line['e2e'] = e2e
    if ns is not None:
        nb = block(ns_plan, ns)
        nroof, nint = rooflines(ns_plan, ns)
        nb.update(config=ns_plan.config, roofline=nroof, int_roofline=nint)
        if ns_cpu is not None:
            nb['cpu_baseline'] = ns_cpu
            nb['x_cpu_resident'] = nb['value'] / ns_cpu['value']
        line['north_star'] = nb
    if cpu is not None:
        line['cpu_baseline'] = cpu
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_e2e(args, plan, shard, lanes, sync_all, rank, world):
    """The same metric through the reference-facing API with HOST buffers.  Input of a step: this
    rank's part of the genome as a packed image (0.375 B/bp, the N4 cache format) in pinned host
    memory — what a resident service keeps, as the reference keeps SeqRecords in RAM.  A step =
    dp_template_load_mem (H2D of the planes) + SearchEngine.find_all over every record and every
    primer (one device pass) + the per (record, primer) result lists on the host."""
    import torch
    import torch.distributed as dist
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.SeqUtils import Seq, SeqRecord
    from degenprimer_b200.SeqView import ResidentRecord
    from degenprimer_b200.WorkCounter import WorkCounter
    from degenprimer_b200 import _native as nat
    n_img = shard.template.image_bytes
    image = torch.empty(n_img, dtype=torch.uint8).pin_memory()
    shard.template.save_image(image.data_ptr(), n_img)
    # an unverified load from pinned memory moves the tables and two of the three planes: the exception
    # plane is rebuilt on the device from the tables (dp_exc_records_kernel / dp_exc_runs_kernel)
    head = np.frombuffer(image.numpy()[:64].tobytes(), dtype=nat.PACK_HEADER)[0]
    n_h2d = n_img - 4 * int(head['n_words_alloc'])
    gprimers = []
    for i, p in enumerate(plan.primers):
        g = Primer(SeqRecord(Seq(p), id='primer%d' % i), 1.5e-6)
        g.generate_components()
        gprimers.append(g)
    # two workers, each with its own context and engine: while one worker's planes are on their
    # way over PCIe the other one's are being scanned (ctypes drops the GIL)
    n_workers = min(2, len(lanes))
    engines = [SearchEngine(jit=not args.no_jit) for _ in range(n_workers)]
    d2h_last = [0] * n_workers
    sites_last = [0] * n_workers

    def e2e_step(w):
        c, se = lanes[w], engines[w]
        t = c.template_load_image(image.data_ptr(), n_img, verify=False)
        recs = [ResidentRecord(t, r) for r in range(t.n_records)]
        # at N > 1 a rank's template holds pieces of records; the pieces are searched as they are
        res = se.find_all(WorkCounter(), recs, gprimers, plan.k)
        d2h_last[w] = 16 * int(se.last_stats['scan']['hits']) + 24 * int(se.last_stats['dg']['dimers_kept'])
        sites_last[w] = sum(len(pair[0]) + len(pair[1]) for pairs in res.values() for pair in pairs if pair)
        t.close()

    def e2e_run(n_steps):
        errors = []

        def work(w):
            try:
                for _ in range(w, n_steps, n_workers):
                    e2e_step(w)
            except BaseException as e:          # noqa: surfaced on the main thread below
                errors.append(e)
        for w in range(1, n_workers):
            helpers.run(w - 1, work, w)
        work(0)
        helpers.wait(n_workers - 1)
        if errors:
            raise errors[0]

    helpers = HostThreads(n_workers - 1)           # long-lived, as in run_ours
    for w in range(n_workers):
        e2e_step(w)
    e2e_run(2 * n_workers)
    sync_all()
    steps = max(4, min(args.steps, 10))
    t0 = time.perf_counter()
    e2e_run(steps)
    sync_all()
    elapsed = time.perf_counter() - t0
    if world > 1:
        tt = torch.tensor([elapsed], dtype=torch.float64, device='cuda')
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        elapsed = float(tt.item())
    helpers.close()
    return {'value': plan.bp * plan.n_var / (elapsed / steps), 'unit': UNIT, 'h2d_bytes_per_step': int(n_h2d), 'image_bytes': int(n_img),
            'd2h_bytes_per_step': int(d2h_last[0]), 'ms_per_step': elapsed / steps * 1e3, 'steps': steps,
            'workers': n_workers, 'sites_per_step': int(sites_last[0]),
            'path': 'packed genome image (0.375 B/bp) in pinned host memory -> dp_template_load_mem (H2D of the tables and '
                    'the two base planes, 0.25 B/bp; the exception plane is rebuilt on the device) -> '
                    'SearchEngine.find_all (all primers x all records, one dp_find) -> per (record, primer) '
                    '(fwd_results, rev_results) on the host, as SearchEngine.find puts them on its shelf; Duplex '
                    'objects are built when a site is looked into, the temp files of `find` are not written '
                    '(neither arm writes them)'}


def main():
    args = parse_args()
    # ONE JSON line on stdout: libraries that chat on fd 1 (NCCL prints its version there on
    # some boxes) are sent to stderr for the duration of the run
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, 'w')
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)
    sys.stdout.flush()


if __name__ == '__main__':
    main()
