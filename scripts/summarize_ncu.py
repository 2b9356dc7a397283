#!/usr/bin/env python
"""Turn ncu outputs into the small text summaries kept under profiles/.

  summarize_ncu.py launches <launches.csv>       per-kernel totals and shares
  summarize_ncu.py full <file.ncu-rep> [cells]   key metrics, stalls, SASS mix
"""
import collections
import csv
import io
import subprocess
import sys


def launches(path):
    rows = [r for r in csv.reader(open(path, errors='replace')) if len(r) > 5]
    hdr = rows[0]
    iK, iV, iU = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
    tot = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[iV].replace(',', ''))
        except ValueError:
            continue
        v *= {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(r[iU], 1.0)     # -> ms
        name = r[iK].split('(')[0]
        t = tot.setdefault(name, [0, 0.0])
        t[0] += 1
        t[1] += v
    total = sum(t[1] for t in tot.values())
    print('# per-launch device time, gpu__time_duration.sum, --clock-control none (serialised, cold cache)')
    print('%-46s %8s %12s %10s %7s' % ('kernel', 'launches', 'total ms', 'mean ms', 'share'))
    for name, (n, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print('%-46s %8d %12.3f %10.4f %6.1f%%' % (name[:46], n, ms, ms / n, 100 * ms / total))
    print('%-46s %8d %12.3f' % ('TOTAL', sum(t[0] for t in tot.values()), total))


WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__warps_active.avg.per_cycle_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum',
        'launch__grid_size', 'launch__block_size', 'local_load', 'local_store']


def full(path, cells=None):
    raw = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    iN = hdr.index('Kernel Name')
    for r in rows[2:]:
        print('## launch:', r[iN])
        vals = dict(zip(hdr, r))
        for w in WANT:
            for h in hdr:
                if h == w or (w.startswith('local') and w in h and h.endswith('.sum')):
                    print('  %-62s %s %s' % (h, vals[h], units[hdr.index(h)]))
        try:
            rd = float(vals['dram__bytes_read.sum']) * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1}[units[hdr.index('dram__bytes_read.sum')]]
            wr = float(vals['dram__bytes_write.sum']) * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1}[units[hdr.index('dram__bytes_write.sum')]]
            print('  dram bytes per launch                                          %.4e' % (rd + wr))
            if cells:
                print('  dram bytes per cell                                            %.2f' % ((rd + wr) / cells))
        except Exception as e:      # noqa: BLE001
            print('  (traffic: %s)' % e)
        stalls = sorted(((float(vals[h]), h) for h in hdr if h.startswith('smsp__pcsamp_warps_issue_stalled_')
                         and not h.endswith('_not_issued')), reverse=True)
        tot = sum(v for v, _ in stalls) or 1.0
        print('  warp stall samples:')
        for v, h in stalls[:8]:
            print('    %-28s %5.1f%%' % (h.replace('smsp__pcsamp_warps_issue_stalled_', ''), 100 * v / tot))
    src = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv', '--print-source', 'sass'],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = rows[1]
    iS, iI = hdr.index('Source'), hdr.index('Instructions Executed')
    cnt, tot = collections.Counter(), 0
    for r in rows[2:]:
        if r and r[0] == 'Kernel Name':
            break
        if len(r) != len(hdr) or r[0] == 'Address':
            continue
        toks = r[iS].split()
        op = (toks[1] if toks[0].startswith('@') else toks[0]).split('.')[0]
        cnt[op] += int(r[iI])
        tot += int(r[iI])
    print('## SASS mix of the first launch: %d warp instructions' % tot
          + (' = %.1f thread instructions per cell' % (tot * 32.0 / cells) if cells else ''))
    for op, c in cnt.most_common(16):
        print('  %-8s %5.1f%%' % (op, 100.0 * c / tot) + ('  %6.1f /cell' % (c * 32.0 / cells) if cells else ''))


if __name__ == '__main__':
    if sys.argv[1] == 'launches':
        launches(sys.argv[2])
    else:
        full(sys.argv[2], float(sys.argv[3]) if len(sys.argv) > 3 else None)
