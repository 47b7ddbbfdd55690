"""Per-source-line executed instructions and stall samples from an ncu report (needs -lineinfo and --import-source on):
  python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top]
Lines of inlined helpers are reported under the file they live in."""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 60
    text = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                          capture_output=True, text=True).stdout
    fname, hdr = None, None
    inst, samp, src = collections.Counter(), collections.Counter(), {}
    stall = collections.defaultdict(collections.Counter)
    for r in csv.reader(io.StringIO(text)):
        if not r:
            continue
        if r[0] == 'File Path':
            fname = r[1].split('/')[-1]
            continue
        if r[0] == 'Line No':
            hdr = {h: i for i, h in enumerate(r)}
            stalls = [(h, i) for i, h in enumerate(r) if h.startswith('stall_') and 'Not Issued' not in h]
            continue
        if hdr is None or not r[0].isdigit():
            continue
        key = (fname, int(r[0]))
        src[key] = r[1].strip()
        num = lambda v: int(v) if v.isdigit() else 0
        inst[key] += num(r[hdr['Instructions Executed']])
        samp[key] += num(r[hdr['# Samples']])
        for h, i in stalls:
            if r[i].isdigit():
                stall[key][h] += int(r[i])
    ti, ts = sum(inst.values()), sum(samp.values())
    print('total warp instructions %d, samples %d' % (ti, ts))
    allst = collections.Counter()
    for k in stall:
        allst.update(stall[k])
    print('stalls:', ', '.join('%s %.1f%%' % (h[6:], 100.0 * n / max(ts, 1)) for h, n in allst.most_common(8)))
    print('--- by instructions')
    for k, n in inst.most_common(top):
        print('%5.2f%% inst %5.2f%% samp  %s:%d  %s' % (100.0 * n / ti, 100.0 * samp[k] / ts, k[0], k[1], src[k][:90]))
    print('--- by samples')
    for k, n in samp.most_common(top):
        worst = ', '.join('%s %d' % (h[6:], c) for h, c in stall[k].most_common(2))
        print('%5.2f%% samp %5.2f%% inst  %s:%d  %s   [%s]' % (100.0 * n / ts, 100.0 * inst[k] / ti, k[0], k[1], src[k][:80], worst))


main()
