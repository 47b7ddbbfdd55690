"""Summarise ncu outputs into the small text files committed under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches.csv  > profiles/rNN_launches.txt
  python tools/ncu_summary.py report   gpurun_out/prof.ncu-rep  > profiles/rNN_kernel.txt
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = ('gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__waves_per_multiprocessor', 'smsp__inst_executed.sum',
        'sm__inst_executed.avg.per_cycle_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__cycles_elapsed.max', 'smsp__cycles_active.avg', 'dram__cycles_active.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed_pipe_lsu.sum',
        'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_alu.sum', 'smsp__warp_issue_stalled',
        'sm__pipe_tensor_cycles_active')


def short(name):
    name = re.sub(r'^void ', '', name)
    name = re.sub(r'\(anonymous namespace\)::', '', name)
    return re.sub(r'\(.*$', '', name)[:90]


def launches(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(io.StringIO(''.join(lines))):
        if r.get('Metric Name') == 'gpu__time_duration.sum':
            v = float(r['Metric Value'].replace(',', ''))
            if r.get('Metric Unit') in ('us', 'usecond'):
                v *= 1e3
            rows.append((short(r['Kernel Name']), v, r['Grid Size'], r['Block Size']))
    agg = collections.OrderedDict()
    for k, v, g, b in rows:
        a = agg.setdefault(k, [0, 0.0, g, b])
        a[0] += 1
        a[1] += v
    total = sum(a[1] for a in agg.values())
    print('# %d launches, %.1f us total (ncu per-launch times: cold-cache, serialised -> compare SHARES)' % (len(rows), total / 1e3))
    print('%-92s %6s %10s %10s %7s  %s' % ('kernel', 'count', 'total_us', 'mean_us', 'share', 'grid x block (last)'))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print('%-92s %6d %10.1f %10.2f %6.1f%%  %s x %s' % (k, a[0], a[1] / 1e3, a[1] / a[0] / 1e3, 100 * a[1] / total, a[2], a[3]))


def report(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print('## %s  grid %s block %s' % (short(d.get('Kernel Name', '?')), d.get('Grid Size'), d.get('Block Size')))
        for h, u, v in zip(hdr, units, r):
            if any(h.startswith(k) for k in KEYS):
                print('  %-75s %16s %s' % (h, v, u))


if __name__ == '__main__':
    {'launches': launches, 'report': report}[sys.argv[1]](sys.argv[2])
