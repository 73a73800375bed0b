#!/usr/bin/env python3
"""profiles/kernel_counters.json from one `ncu --set full` capture that holds one config-4 step followed by one
north-star step (tools/ncu_full.sh <tag> 21 6):   python tools/refresh_counters.py <raw.csv> <source text>"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BP = 3099999988.0          # synthetic.GENOME_3G


def main(path, source):
    rows = list(csv.reader(open(path)))
    idx = {h: i for i, h in enumerate(rows[0])}
    units = rows[1]

    def val(r, name):
        v = float(r[idx[name]].replace(',', ''))
        u = units[idx[name]]
        scale = {'Mbyte': 1e6, 'Kbyte': 1e3, 'Gbyte': 1e9, 'byte': 1.0, 'us': 1e-3, 'ms': 1.0, 'ns': 1e-6, 's': 1e3}
        return v * scale.get(u, 1.0)
    scans = [r for r in rows[2:] if r[idx['Kernel Name']].startswith('dp_jit_scan')]
    dups = [r for r in rows[2:] if r[idx['Kernel Name']].startswith('dp_duplex_kernel')]
    if len(scans) < 2:
        raise SystemExit('expected a config-4 and a north-star dp_jit_scan launch in %s' % path)
    out_path = os.path.join(ROOT, 'profiles', 'kernel_counters.json')
    d = json.load(open(out_path))

    def scan_entry(r):
        return dict(warp_inst_per_bp=val(r, 'smsp__inst_executed.sum') / BP,
                    dram_bytes_per_launch=int(val(r, 'dram__bytes_read.sum') + val(r, 'dram__bytes_write.sum')),
                    alu_pipe_pct=val(r, 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active'),
                    issue_active_pct=val(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'),
                    registers=int(val(r, 'launch__registers_per_thread')),
                    ms_under_ncu=val(r, 'gpu__time_duration.sum'), source=source)
    d['config4_20plex_3Gbp']['dp_jit_scan'] = scan_entry(scans[0])
    d['northstar_256fold_3Gbp']['dp_jit_scan'] = scan_entry(scans[1])
    if len(dups) >= 2:
        r = dups[1]
        d['northstar_256fold_3Gbp']['dp_duplex_kernel'] = dict(
            warp_inst_per_launch=val(r, 'smsp__inst_executed.sum'), ms_under_ncu=val(r, 'gpu__time_duration.sum'),
            issue_active_pct=val(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'),
            alu_pipe_pct=val(r, 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active'),
            fp64_pipe_pct=val(r, 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'),
            registers=int(val(r, 'launch__registers_per_thread')), source=source)
    json.dump(d, open(out_path, 'w'), indent=1)
    print(json.dumps(d['config4_20plex_3Gbp']['dp_jit_scan']))


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2])
