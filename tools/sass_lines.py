"""Attribute ncu per-SASS-instruction counters to source lines.

  python tools/sass_lines.py gpurun_out/prof.ncu-rep fast_kernelILi8 [poses] > profiles/rNN_lines.txt

ncu's CSV source page prints counters per SASS instruction only; nvdisasm -g gives the line of every instruction of
the same cubin (extracted from the in-tree .so).  The two are joined by instruction order.  For inlined device
functions the line is attributed to the call site inside the kernel when nvdisasm prints the inline chain.
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'pedestrians_video_2_carla_b200', 'libp2c_b200.so')


def disasm(kernel_pat):
    tmp = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', 'all', LIB], cwd=tmp, stdout=subprocess.DEVNULL, check=True)
    for f in sorted(os.listdir(tmp)):
        out = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(tmp, f)], stdout=subprocess.PIPE, text=True).stdout
        if kernel_pat in out:
            return out
    raise SystemExit('kernel %s not found' % kernel_pat)


def instr_lines(text, kernel_pat):
    lines, cur, on, chain = [], None, False, []
    for ln in text.splitlines():
        if ln.startswith('//---') and '.text.' in ln:
            on = kernel_pat in ln
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            chain = [cur]
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+) inlined at', ln)
        if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', ln):
            op = re.sub(r'/\*[0-9a-f]+\*/', '', ln).strip()
            lines.append((cur, op))
    return lines


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    poses = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'],
                         stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = next(r for r in rows if r and r[0] == 'Address')
    body = [r for r in rows[rows.index(hdr) + 1:] if len(r) == len(hdr)]
    ie, sm = hdr.index('Instructions Executed'), hdr.index('# Samples')
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    src = instr_lines(disasm(pat), pat)
    if len(src) != len(body):
        print('# WARNING: %d SASS rows in the report vs %d in the cubin (rebuilt since the capture?)' % (len(body), len(src)))
    src_text = {}
    for fn in set(s[0][0] for s in src if s[0]):
        p = os.path.join(ROOT, 'pedestrians_video_2_carla_b200', 'csrc', fn)
        if os.path.exists(p):
            src_text[fn] = open(p).read().splitlines()
    agg = collections.OrderedDict()
    tot_i = tot_s = 0
    stalls = collections.Counter()
    for (loc, op), r in zip(src, body):
        n, s = int(r[ie] or 0), int(r[sm] or 0)
        a = agg.setdefault(loc, [0, 0, collections.Counter()])
        a[0] += n
        a[1] += s
        a[2][op.split()[0].lstrip('@!P0123456789 ').split('.')[0] if op else '?'] += n
        tot_i += n
        tot_s += s
        for i in stall_cols:
            if r[i] and r[i] != '0':
                stalls[hdr[i]] += int(r[i])
    print('# %s: %d warp instructions executed (%.1f per pose), %d stall samples' % (pat, tot_i, tot_i / poses, tot_s))
    print('# stall reasons: ' + ', '.join('%s %.1f%%' % (k, 100.0 * v / max(tot_s, 1)) for k, v in stalls.most_common(8)))
    print('%9s %6s %6s  %-22s %s' % ('inst/pose', 'inst%', 'smpl%', 'where', 'source'))
    for loc, (n, s, ops) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:60]:
        text = ''
        if loc and loc[0] in src_text and loc[1] <= len(src_text[loc[0]]):
            text = src_text[loc[0]][loc[1] - 1].strip()[:90]
        print('%9.1f %5.1f%% %5.1f%%  %-22s %s' % (n / poses, 100.0 * n / max(tot_i, 1), 100.0 * s / max(tot_s, 1),
                                                   '%s:%d' % loc if loc else '?', text))


if __name__ == '__main__':
    main()
