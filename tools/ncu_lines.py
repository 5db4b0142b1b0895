#!/usr/bin/env python3
"""Per CUDA source line: warp instructions executed and stall samples of one kernel of an .ncu-rep
(ncu -i REP --page source --csv --print-source cuda,sass).  usage: tools/ncu_lines.py REP KERNEL_REGEX [min_pct]"""
import csv, io, subprocess, sys, collections

def main():
    rep, kern = sys.argv[1], sys.argv[2]
    min_pct = float(sys.argv[3]) if len(sys.argv) > 3 and sys.argv[3][0] != '-' else 0.4
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '--kernel-name', 'regex:' + kern],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    inst, samp, text, thr = collections.Counter(), collections.Counter(), {}, collections.Counter()
    fname, hdr, seen_kernel = '?', None, 0
    addr_inst, addr_samp, addr_thr, addr_outer = {}, {}, {}, {}
    for r in rows:
        if not r: continue
        if r[0] == 'Kernel Name':
            seen_kernel += 1
            if seen_kernel > 1: break
            continue
        if r[0] == 'File Name': fname = r[1].split('/')[-1]; continue
        if r[0] == 'Line No': hdr = {h: i for i, h in enumerate(r)}; ia = r.index('Address'); continue
        if hdr is None or len(r) < 8: continue
        if r[0]:  # a source line row
            cur = (fname, int(r[0])); text[cur] = r[1]
        if r[ia].startswith('0x'):
            try:
                inst[cur] += int(float(r[hdr['Instructions Executed']] or 0)); samp[cur] += int(float(r[hdr['# Samples']] or 0))
                thr[cur] += int(float(r[hdr['Thread Instructions Executed']] or 0))
                addr_inst[r[ia]] = int(float(r[hdr['Instructions Executed']] or 0)); addr_samp[r[ia]] = int(float(r[hdr['# Samples']] or 0))
                addr_thr[r[ia]] = int(float(r[hdr['Thread Instructions Executed']] or 0))
                if r[ia] not in addr_outer or cur[1] > addr_outer[r[ia]][1]: addr_outer[r[ia]] = cur
            except ValueError:
                pass
    ti, ts = sum(addr_inst.values()) or 1, sum(addr_samp.values()) or 1
    if '--outer' in sys.argv:  # every instruction counted once, under the outermost (highest-numbered) line that lists it
        inst, samp, thr = collections.Counter(), collections.Counter(), collections.Counter()
        for a, k in addr_outer.items():
            inst[k] += addr_inst[a]; samp[k] += addr_samp[a]; thr[k] += addr_thr[a]
    print('%s: %d warp instructions, %d samples (a line that calls an inlined function includes the callee: inclusive counts)' % (kern, ti, ts))
    for k in sorted(inst):
        if 100 * inst[k] / ti >= min_pct or 100 * samp[k] / ts >= min_pct:
            print('%-12s %5d  inst %5.2f%%  samp %5.2f%%  lanes %4.1f  %s' % (k[0][:12], k[1], 100 * inst[k] / ti, 100 * samp[k] / ts,
                  thr[k] / max(inst[k], 1), text[k].strip()[:110]))

if __name__ == '__main__':
    main()
