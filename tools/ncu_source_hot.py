#!/usr/bin/env python3
"""Top source lines of a kernel in an ncu report by executed warp instructions and stall samples.
Usage: ncu_source_hot.py report.ncu-rep kernel_regex [N]"""
import csv, subprocess, sys
def main(rep, kern, top=30):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kern, '--print-source', 'cuda,sass'],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    cur, hdr, agg = None, None, []
    for r in rows:
        if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
        if r and r[0] == 'Line No': hdr = r; continue
        if hdr and len(r) > 8 and r[0] not in ('', 'Line No'):
            try:
                agg.append((cur, int(r[0]), r[1].strip(), int(r[hdr.index('Instructions Executed')]), int(r[hdr.index('# Samples')])))
            except ValueError:
                pass
    ti = sum(a[3] for a in agg) or 1; ts = sum(a[4] for a in agg) or 1
    print('total warp inst %d, samples %d' % (ti, ts))
    for a in sorted(agg, key=lambda a: -a[4])[:top]:
        print('%5.1f%% smp %5.1f%% inst  %s:%d  %s' % (100.0 * a[4] / ts, 100.0 * a[3] / ti, a[0], a[1], a[2][:110]))
if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 30)
