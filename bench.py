#!/usr/bin/env python3
"""bench.py — the hot path's headline metric on B200 (contract: task brief ④).

metric : template bp x primer-variants scanned per second (BASELINE.json), with the
         dG stage run on every hit inside the same step.
step   : one pass of the hot path over one resident template: dp_scan (all
         primers, both strands) + dp_duplex_dg (every hit x every variant).
N = 1  : BASELINE.json configs[1] — TP53 pair vs a synthetic 83,257,441 bp
         chr17-sized template, k = 1 (`config2_tp53_chr17`).
N > 1  : weak scaling — rank g holds segment g (83 Mbp + (L-1) halo) of an
         N x 83 Mbp genome, scans it alone, and hit records are all-gathered
         over NVLink peer memory every step (no collective in the scan itself).

`python bench.py --gpus N --steps K --warmup W [--impl reference]`
prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'template_bp_x_primer_variants_scanned_per_s'
UNIT = 'bp*variants/s'
WORKLOAD = 'config2_tp53_chr17'
COND = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0)     # reference README.md:72-96
L2_BYTES = 126 * 1024 * 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default=WORKLOAD)
    ap.add_argument('--scale', type=float, default=1.0, help='shrink the template (testing only)')
    ap.add_argument('--algo', type=int, default=0, help='0 auto, 1 popcount, 2 bit-slice')
    ap.add_argument('--cpu-sample-bp', type=int, default=83_257_441,
                    help='template prefix the cpu_baseline leg of the GPU arm times (about 12 s on 16 cores)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-jit', action='store_true', help='generic scan kernel instead of the run-time specialised one')
    ap.add_argument('--host-threads', type=int, default=2, help='host threads that queue and collect the calls')
    ap.add_argument('--in-flight', type=int, default=4,
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


# ---- clocks -------------------------------------------------------------------------
class ClockSampler(object):
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md) through
    NVML in a background thread (every 2 ms; the nvidia-smi CLI takes longer to start than the
    timed region lasts)."""
    REASONS = (('hw_slowdown', 0x8), ('hw_thermal_slowdown', 0x40), ('sw_thermal_slowdown', 0x20),
               ('sw_power_cap', 0x4))

    def __init__(self, device):
        self.device = device
        self.thread = None
        self.samples, self.maxes, self.bits = [], [], 0
        self._stop = False

    def start(self):
        import threading
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
                time.sleep(0.002)
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


# ---- CPU arm: the reference algorithm (oracle O1) on the host cores -------------------------
def cpu_reference_pass(text, primers, k, workers):
    """One pass of the reference's path (SearchEngine.find per primer: FFT scan
    over multiprocessing slices, then Duplex compilation of every hit x variant).
    Only bench.py's cpu_baseline / --impl reference legs call into oracle/."""
    from oracle import conditions as ocond
    from oracle import search_o1 as o1
    from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord
    ocond.PCR_P.set(COND)
    hits = dimers = 0
    for i, master in enumerate(primers):
        m = o1.find_mp(text, master, k, workers=workers)
        p = OPrimer(OSeqRecord(master, id='p%d' % i), 1.5e-6)
        res = o1.compile_duplexes(m[0], m[1], p, m[2], m[3], m[4], m[5], mp_order=True)
        hits += len(m[4]) + len(m[5])
        if res:
            dimers += sum(len(d.dimers) for lst in res for _, dups in lst for d in dups)
    return hits, dimers


def make_cpu_sample(wl, sample_bp):
    """A bounded prefix of the workload's template that still contains planted sites."""
    rec = wl['records'][0]
    n = min(len(rec), sample_bp)
    arr = rec[:n].copy()
    from degenprimer_b200.synthetic import plant_pairs
    plant_pairs(np.random.default_rng(99), arr, wl['primers'], wl['k'], 8)
    return arr.tobytes().decode()


def cpu_baseline(wl, sample_bp, repeats=1):
    cores = os.cpu_count() or 1
    text = make_cpu_sample(wl, sample_bp)
    n_var = sum(len(v) for v in variants_of(wl['primers']))
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        hits, dimers = cpu_reference_pass(text, wl['primers'], wl['k'], cores)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return dict(value=len(text) * n_var / best, unit=UNIT, cores=cores, kind='port',
                sample='first %d bp of %s x %d primers, scan (scipy.fftpack, %d fork workers) + dG of %d hits; %.2f s'
                       % (len(text), wl['name'], len(wl['primers']), cores, hits, best),
                seconds=best, hits=hits, dimers=dimers)


def run_reference(args):
    """--impl reference: the reference's own CPU algorithm for this path.  The
    reference itself is Python 2.7 + Biopython and cannot run in this image
    (no python2, no Bio), so its Python-3 restatement (oracle O1)
    is what is timed; `kind` says so."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from degenprimer_b200.synthetic import make_workload
    wl = make_workload(args.workload, scale=args.scale)
    cores = os.cpu_count() or 1
    n_var = sum(len(v) for v in variants_of(wl['primers']))
    # the per-step sample is sized so that the whole run (K timed steps) takes about 2.5 minutes
    # on this host, between 1 Mbp and the whole template; throughput is per bp
    probe = make_cpu_sample(wl, 2_000_000)
    cpu_reference_pass(probe[:500_000], wl['primers'], wl['k'], cores)        # fork pool, imports
    t0 = time.perf_counter()
    cpu_reference_pass(probe, wl['primers'], wl['k'], cores)
    s_per_bp = (time.perf_counter() - t0) / max(1, len(probe))
    budget = min(15.0, max(0.25, 150.0 / max(1, args.steps)))
    sample_bp = int(min(wl['bp'], max(1_000_000, budget / s_per_bp)))
    text = make_cpu_sample(wl, sample_bp)
    for _ in range(args.warmup):
        cpu_reference_pass(text[:min(len(text), 1_000_000)], wl['primers'], wl['k'], cores)
    t0 = time.perf_counter()
    hits = 0
    for _ in range(args.steps):
        hits, _d = cpu_reference_pass(text, wl['primers'], wl['k'], cores)
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    value = len(text) * n_var / dt
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'complex128 FFT (reference) / u32 bitwise (ours)',
        'data': 'synthetic',
        'config': {'workload': args.workload, 'primers': wl['primers'], 'max_mismatches': wl['k'],
                   'step_sample_bp': len(text), 'note': 'bounded sample per step; throughput is per bp so it is size independent'},
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                         'sample': '%d bp x %d primers per step, %d hits' % (len(text), len(wl['primers']), hits)},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line))


def bind_near_gpu(local):
    """Pins this rank to the host cores NVML lists as local to its GPU, so that the pinned
    staging memory of the e2e leg is first touched on the GPU's own NUMA node (8 ranks copy
    83 MB per step each)."""
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


# ---- GPU arm -----------------------------------------------------------------------------------
def run_ours(args):
    rank, world, local = dist_env()
    if world > 1:
        bind_near_gpu(local)
    if world != args.gpus and world > 1:
        raise SystemExit('WORLD_SIZE=%d but --gpus %d' % (world, args.gpus))
    from degenprimer_b200.synthetic import make_workload
    wl = make_workload(args.workload, scale=args.scale, seed_offset=1000 * rank)
    primers, k = wl['primers'], wl['k']
    var_lists = variants_of(primers)
    n_var = sum(len(v) for v in var_lists)
    bp = wl['bp']

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(wl, args.cpu_sample_bp)      # before CUDA is initialised (fork pool)

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    os.environ['DEGENPRIMER_B200_DEVICE'] = str(local)
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import _native as nat
    from degenprimer_b200.distributed import all_gather_records, sort_hits
    tdf.PCR_P.set(COND)
    ctx = tdf._ctx()
    info = ctx.device_info()
    cond = tdf.conditions()

    # inputs: pinned host ASCII (for e2e) and rotating resident copies (> L2) for the resident arm
    host = [torch.from_numpy(r).pin_memory() for r in wl['records']]
    srcs = [(h.data_ptr(), h.numel()) for h in host]
    one = ctx.template_from_sequences(srcs)
    plane_bytes = one.device_bytes
    n_copies = max(2, int(np.ceil(1.5 * L2_BYTES / max(1, plane_bytes))))
    n_copies = min(n_copies, 64)
    copies = [one] + [ctx.template_from_sequences(srcs) for _ in range(n_copies - 1)]

    # Lanes: contexts (own stream, own result buffers) on this GPU.  The resident arm keeps one
    # call in flight per lane, so the host work of a call (queueing ~30 us, result conversion)
    # overlaps the kernels of the other lane's call; templates are shared between the lanes.
    n_lanes = max(1, args.in_flight)
    lanes = [ctx] + [tdf.new_context(local) for _ in range(n_lanes - 1)]
    # the primer side, laid out once; a throughput loop over one query is what the run-time
    # specialised scan kernel is for (compiled here, ~0.3 s, outside every timed region)
    queries = [c.prepare(primers, k, var_lists, jit=not args.no_jit) for c in lanes]
    exchs = [None] * n_lanes
    if world > 1:
        # the hit gather is the epilogue of the publish kernel: each rank stores its slot into
        # every peer's receive buffer over NVLink (CUDA IPC peer memory), no collective launch
        from degenprimer_b200.distributed import HitExchange
        exchs = [HitExchange(c, max_hits=1024) for c in lanes]

    def gathered(lane, age, own):
        # hits of every rank for a call of this lane (rank g's segment = record g of the genome).
        # The publish kernel of the lane's next call moves a call's slots to pinned host memory,
        # so inside the loop the call two back is collected while the newest one runs (age 2:
        # conversion and sort overlap the kernels); the last two calls of a lane are collected
        # after the loop (ages 1 and 0) — every gather completes inside the timed region.
        allh = exchs[lane].collect_hits(age)
        return allh if allh is not None else sort_hits(all_gather_records(own))

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

    def drive(my, n_steps, first_call, acc):
        """n_steps steps on the lanes `my` (indices), one call in flight per lane, from one
        host thread.  A step = one pass of the hot path: scan of every primer over both
        strands + dG of every (hit, variant), one stream submission (dp_find), results back
        on the host."""
        F = len(my)
        made = [0] * F                          # calls of this region queued on a lane
        own = [[None, None] for _ in range(F)]  # a lane's own hits of its last two calls

        def begin(i):
            j = i % F
            lanes[my[j]].find_begin(copies[(first_call + i) % n_copies], queries[my[j]], cond, algo=args.algo)
            made[j] += 1

        for i in range(min(F - 1, n_steps)):
            begin(i)
        for i in range(n_steps):
            if i + F - 1 < n_steps:
                begin(i + F - 1)
            j = i % F
            lane = my[j]
            if exchs[lane] is not None and made[j] >= 3:
                acc['hits'] += len(gathered(lane, 2, own[j][0]))
            hits, dimers, st, dst = lanes[lane].find_end()
            own[j] = [own[j][1], hits]
            if exchs[lane] is None:
                acc['hits'] += len(hits)
            acc['ms_main'] += st['ms_main']
            acc['ms_scan'] += st['ms_scan']
            acc['ms_dg'] += dst['ms_kernel']
            acc['launches'] += st['launches'] + dst['launches']
            acc['dimers'] += len(dimers)
            acc['cand'] += st['candidates']
            acc['windows'] = st['windows']
            acc['algo'] = st['algo']
        for j in range(F):
            lane = my[j]
            if exchs[lane] is not None:
                if made[j] >= 2:
                    acc['hits'] += len(gathered(lane, 1, own[j][0]))
                if made[j] >= 1:
                    acc['hits'] += len(gathered(lane, 0, own[j][1]))

    def timed_region(K, F, T, first_call):
        """K steps with F calls in flight (one per lane), queued and collected by T host threads
        (F = T = 1: strictly one call after the other on the main thread).  Step i of the
        region belongs to thread i % T, so every rank makes the same calls on the same lane."""
        import threading
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        T = max(1, min(T, F))
        accs = [dict(ms_main=0.0, ms_scan=0.0, ms_dg=0.0, launches=0, hits=0, dimers=0, cand=0, windows=0, algo=0)
                for _ in range(T)]
        shares = [K // T + (1 if t < K % T else 0) for t in range(T)]
        firsts = [first_call + sum(shares[:t]) for t in range(T)]
        groups = [list(range(t, F, T)) for t in range(T)]
        for c in lanes:
            c.set_timing(F == 1)               # the ms_* stats are only read in the sequential region
        errors = []

        def work(t):
            try:
                drive(groups[t], shares[t], firsts[t], accs[t])
            except BaseException as e:          # noqa: surfaced on the main thread below
                errors.append(e)

        sync_all()
        t0 = time.perf_counter()
        ev0.record(kstream)
        threads = [threading.Thread(target=work, args=(t,)) for t in range(1, T)]
        for th in threads:
            th.start()
        work(0)
        for th in threads:
            th.join()
        ev1.record(kstream)
        if errors:
            raise errors[0]
        sync_all()
        wall = time.perf_counter() - t0
        elapsed = ev0.elapsed_time(ev1) * 1e-3        # device clock, first launch -> last result on the host
        if world > 1:
            tt = torch.tensor([elapsed, wall], dtype=torch.float64, device='cuda')
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            elapsed, wall = float(tt[0].item()), float(tt[1].item())
        acc = accs[0]
        for o in accs[1:]:
            for key in ('ms_main', 'ms_scan', 'ms_dg', 'launches', 'hits', 'dimers', 'cand'):
                acc[key] += o[key]
        return elapsed, wall, acc

    K = max(1, args.steps)
    W = max(3, args.warmup)
    n_threads = max(1, min(args.host_threads, n_lanes))
    timed_region(max(W, 3 * n_lanes), n_lanes, n_threads, 0)      # warm-up, untimed (every lane, every thread)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # (1) the headline: K steps, n_lanes calls in flight
    elapsed, wall, acc = timed_region(K, n_lanes, n_threads, W)
    ms_step = elapsed / K * 1e3
    value = world * bp * n_var / (elapsed / K)
    # (2) the same K steps strictly one after the other: per-call latency, and kernel durations
    #     that are not stretched by the other lane's kernels (the roofline below uses these)
    seq_elapsed, seq_wall, seq = timed_region(K, 1, 1, W + K)
    for ex in exchs:
        if ex is not None:
            ex.close()          # the e2e leg below runs the ranks independently

    # ---- e2e: host ASCII in pinned memory -> reference-facing calls -> host results -------------
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.SeqUtils import Seq, SeqRecord
    from degenprimer_b200.SeqView import ResidentRecord
    from degenprimer_b200.WorkCounter import WorkCounter
    gprimers = []
    for i, p in enumerate(primers):
        g = Primer(SeqRecord(Seq(p), id='primer%d' % i), 1.5e-6)
        g.generate_components()
        gprimers.append(g)
    # two worker threads, each with its own context and SearchEngine: while one worker's
    # template is on its way over PCIe the other one's is being scanned (ctypes drops the GIL);
    # the leg is PCIe bound (83 MB per step), more workers do not help
    import threading
    n_workers = min(2, n_lanes)
    engines = [SearchEngine() for _ in range(n_workers)]
    d2h_last = [0] * n_workers

    def e2e_step(lane):
        c, se = lanes[lane], engines[lane]
        t = c.template_from_sequences(srcs)                   # H2D of the ASCII bytes + pack kernel
        d2h = 0
        for r in range(t.n_records):
            rec = ResidentRecord(t, r)
            for g in gprimers:
                se.find_duplexes(WorkCounter(), rec, g, k)    # SearchEngine.find without the shelf file
                d2h += 16 * int(se.last_stats['scan']['hits']) + 24 * int(se.last_stats['dg']['dimers_kept'])
        t.close()
        d2h_last[lane] = d2h

    def e2e_run(n_steps):
        errors = []

        def work(lane):
            try:
                for i in range(lane, n_steps, n_workers):
                    e2e_step(lane)
            except BaseException as e:          # noqa: surfaced on the main thread below
                errors.append(e)
        threads = [threading.Thread(target=work, args=(lane,)) for lane in range(1, n_workers)]
        for th in threads:
            th.start()
        work(0)
        for th in threads:
            th.join()
        if errors:
            raise errors[0]

    for lane in range(n_workers):
        e2e_step(lane)
    e2e_run(2 * n_workers)
    sync_all()
    e2e_steps = max(4, min(args.steps, 10 * n_workers))
    t0 = time.perf_counter()
    e2e_run(e2e_steps)
    sync_all()
    e2e_elapsed = time.perf_counter() - t0
    d2h = d2h_last[0]
    if world > 1:
        tt = torch.tensor([e2e_elapsed], dtype=torch.float64, device='cuda')
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_elapsed = float(tt.item())
    e2e_value = world * bp * n_var / (e2e_elapsed / e2e_steps)
    clocks = sampler.stop() if rank == 0 else None

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
    ms_main = seq['ms_main'] / K              # from the sequential region: not stretched by the other lane
    algo_bytes = 0.25 * bp                      # 2 bits per template base, read once
    achieved = algo_bytes / (ms_main * 1e-3) / 1e9 if ms_main > 0 else 0.0
    kernel = {2: 'dp_scan_bitslice_kernel', 3: 'dp_jit_scan'}.get(seq['algo'], 'dp_scan_popc_kernel')
    traffic = None
    try:
        with open(os.path.join(ROOT, 'profiles', 'traffic.json')) as fh:
            entry = json.load(fh).get(args.workload, {}).get(kernel)
        if entry and args.scale == 1.0:
            traffic = entry['bytes_per_launch']
    except Exception:
        pass
    roofline = {'bound': 'hbm', 'achieved': achieved, 'peak': hbm_peak, 'unit': 'GB/s',
                'frac': achieved / hbm_peak, 'traffic': traffic,
                'kernel': kernel,
                'ms_per_launch': ms_main, 'algorithmic_bytes_per_launch': algo_bytes,
                'peak_source': peak_src}
    # integer-pipe view (the path is integer/bitwise work; canonical op counts of SURVEY §8d)
    lop3 = ctx.calibrate('lop3')
    popc = ctx.calibrate('popc')
    lds = ctx.calibrate('lds')
    windows = float(seq['windows'])
    a_win = 16.0
    int_roofline = {
        'bound': 'int32 ALU (LOP3) issue rate, measured by dp_calibrate on this GPU',
        'peak_lop3_lane_ops_per_s': lop3, 'peak_popc_lane_ops_per_s': popc, 'peak_lds32_lane_ops_per_s': lds,
        'windows_per_s': windows / (ms_main * 1e-3) if ms_main > 0 else 0.0,
        'canonical_ops_per_window': a_win,
        'frac_of_canonical_int_roof': (windows * a_win / (ms_main * 1e-3)) / lop3 if ms_main > 0 else 0.0,
        'frac_of_canonical_popc_roof': (windows * 2.0 / (ms_main * 1e-3)) / popc if ms_main > 0 else 0.0,
        'note': 'canonical = 16 int32 ops incl. 2 POPC per window (SURVEY 8d); the bit-sliced kernel '
                'issues far fewer, so fractions above 1 mean "faster than the per-window popcount roof"',
    }
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': ms_step, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'u32 bitwise (scan) + f64 (dG)', 'data': 'synthetic',
        'config': {'workload': args.workload, 'template_bp_per_gpu': bp, 'primers': primers,
                   'variants': n_var, 'max_mismatches': k, 'records': len(wl['records']),
                   'l2': 'rotating %d resident copies of the template planes (%.0f MB > 126 MB L2)'
                         % (n_copies, n_copies * plane_bytes / 1e6),
                   'calls_in_flight': n_lanes, 'host_threads': n_threads,
                   'timing': 'CUDA events around K steps with %d calls in flight (one context each, %d host '
                             'threads; max over ranks); wall clock %.4f ms/step' % (n_lanes, n_threads, wall / K * 1e3),
                   'gpu': info['name'], 'sms': info['sm_count'], 'scale': args.scale},
        'hits_per_step': acc['hits'] / K, 'dimers_per_step': acc['dimers'] / K,
        'candidates_per_step': acc['cand'] / K,
        'ms_scan_kernels': seq['ms_scan'] / K, 'ms_dg_kernels': seq['ms_dg'] / K,
        'dg_hits_per_s': (seq['hits'] / K) / (seq['ms_dg'] / K * 1e-3) if seq['ms_dg'] > 0 else None,
        'gpu_launches': acc['launches'] + K,          # scan kernels + duplex kernel + the publish kernel of every step
        'sequential': {'ms_per_step': seq_elapsed / K * 1e3, 'value': world * bp * n_var / (seq_elapsed / K),
                       'note': 'the same K steps with one call in flight (per-call latency); the kernel '
                               'durations of the roofline are taken here'},
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': int(sum(n for _, n in srcs)),
                'd2h_bytes_per_step': int(d2h), 'ms_per_step': e2e_elapsed / e2e_steps * 1e3,
                'workers': n_workers,
                'path': 'pinned host ASCII -> dp_template_from_records (H2D + pack) -> '
                        'SearchEngine.find_duplexes (= find without the shelf file) per primer -> host arrays'},
        'roofline': roofline, 'int_roofline': int_roofline, 'clocks': clocks,
    }
    if cpu is not None:
        line['cpu_baseline'] = cpu
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


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
