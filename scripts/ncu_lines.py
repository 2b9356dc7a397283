#!/usr/bin/env python
"""Per-source-line instruction and stall-sample counts of the first kernel in an
ncu report (needs -lineinfo and --import-source on):
    python scripts/ncu_lines.py gpurun_out/prof_X.ncu-rep [cells] [top]"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    cells = float(sys.argv[2]) if len(sys.argv) > 2 else None
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                         capture_output=True, text=True).stdout
    rows, fname, hdr, kernels = [], None, None, 0
    for r in csv.reader(io.StringIO(out)):
        if not r:
            continue
        if r[0] == 'File Path':
            fname = r[1].split('/')[-1]
            continue
        if r[0] == 'Function Name':
            continue
        if r[0] == 'Line No':
            hdr = r
            if fname and fname in [x[0] for x in rows[:1]]:
                pass
            continue
        if hdr and len(r) > 8 and r[2] == '-' and r[0].isdigit():
            iI = hdr.index('Instructions Executed')
            iS = hdr.index('# Samples')
            rows.append((fname, int(r[0]), r[1].strip(), int(r[iI] or 0), int(r[iS] or 0)))
    # the report holds every captured launch one after another: keep the first occurrence
    seen, uniq = set(), []
    for r in rows:
        k = (r[0], r[1])
        if k in seen:
            continue
        seen.add(k)
        uniq.append(r)
    tot_i = sum(r[3] for r in uniq)
    tot_s = sum(r[4] for r in uniq)
    print('total warp instructions %d' % tot_i + (' = %.1f thread instr / cell' % (tot_i * 32 / cells) if cells else ''))
    print('%-22s %5s %7s %7s  %s' % ('file', 'line', 'instr%', 'stall%', 'source'))
    for r in sorted(uniq, key=lambda r: -r[3])[:top]:
        print('%-22s %5d %6.1f%% %6.1f%%  %s' % (r[0], r[1], 100.0 * r[3] / tot_i, 100.0 * r[4] / max(tot_s, 1), r[2][:90]))
    print('--- by stall samples')
    for r in sorted(uniq, key=lambda r: -r[4])[:15]:
        print('%-22s %5d %6.1f%% %6.1f%%  %s' % (r[0], r[1], 100.0 * r[3] / tot_i, 100.0 * r[4] / max(tot_s, 1), r[2][:90]))


if __name__ == '__main__':
    main()
