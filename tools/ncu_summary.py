#!/usr/bin/env python3
"""Summarise an .ncu-rep (run here, no GPU needed): key metrics per kernel + per-line hot spots.
usage: tools/ncu_summary.py report.ncu-rep [kernel-regex-for-source-page ...]"""
import collections
import csv
import io
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__grid_size', 'inst_executed', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'sm__cycles_elapsed.max',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']


def raw(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('##', r[hdr.index('Kernel Name')])
        for w in WANT:
            if w in hdr:
                print('  %-62s %s %s' % (w, r[hdr.index(w)], units[hdr.index(w)]))


def source(rep, kernel, top=28):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kernel], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[2:] if len(r) >= len(hdr) - 2 and r[0].startswith('0x')]
    half = len(data) // 2
    if half and data[0][ix['Source']] == data[half][ix['Source']] and data[0][0] == data[half][0]:
        data = data[:half]  # the page lists the kernel twice
    gi = lambda r, k: int(float(r[ix[k]] or 0))
    ti = sum(gi(r, 'Instructions Executed') for r in data)
    ts = sum(gi(r, '# Samples') for r in data) or 1
    print('source page %s: %d SASS lines, %d warp-instructions, %d samples' % (kernel, len(data), ti, ts))
    ops, samp = collections.Counter(), collections.Counter()
    for r in data:
        t = r[ix['Source']].strip().split()
        op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        ops[op] += gi(r, 'Instructions Executed')
        samp[op] += gi(r, '# Samples')
    print('opcode mix (inst%/samples%):', ', '.join('%s %.1f/%.1f' % (o, 100 * c / ti, 100 * samp[o] / ts) for o, c in ops.most_common(22)))
    st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    agg = {h: sum(gi(r, h) for r in data) for h in st}
    print('stalls:', ', '.join('%s %.1f%%' % (k[6:], 100 * v / ts) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:9]))
    for i in sorted(sorted(range(len(data)), key=lambda i: -gi(data[i], '# Samples'))[:top]):
        r = data[i]
        print('  %4d %5.2f%% x%-9d %s' % (i, 100 * gi(r, '# Samples') / ts, gi(r, 'Instructions Executed'), r[ix['Source']].strip()[:80]))


if __name__ == '__main__':
    raw(sys.argv[1])
    for k in sys.argv[2:]:
        source(sys.argv[1], k)
