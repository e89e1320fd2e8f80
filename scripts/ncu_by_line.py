#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` SASS listing of one kernel by CUDA source line.

usage: ncu_by_line.py <source.csv> <nvdisasm -g -c listing> <kernel mangled-name substring> [block]
Joins the i-th SASS instruction of the ncu listing with the i-th instruction of the nvdisasm
listing (same cubin, same order) to get file:line, then prints instruction share, stall-sample
share and average active threads per source line (inlined call sites resolved to the innermost line).
"""
import csv, re, sys, collections

def main():
    src, lst, kname = sys.argv[1:4]
    blk = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    rows = list(csv.reader(open(src)))
    hidx = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
    h = rows[hidx[blk]]
    end = hidx[blk + 1] - 1 if blk + 1 < len(hidx) else len(rows)
    ci, si, ti = h.index('Instructions Executed'), h.index('# Samples'), h.index('Thread Instructions Executed')
    data = [r for r in rows[hidx[blk] + 1:end] if len(r) > ci and r[ci].isdigit()]
    # nvdisasm listing
    lines = open(lst).read().split('\n')
    start = next(i for i, l in enumerate(lines) if l.startswith('.text.') and kname in l)
    loc = None
    locs = []
    for l in lines[start + 1:]:
        if l.startswith('.text.') or l.startswith('.section'):
            if locs: break
        m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
        if m:
            loc = (m.group(1).split('/')[-1], int(m.group(2)))
            continue
        if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', l):
            locs.append(loc)
    if len(locs) != len(data):
        print(f'warning: {len(locs)} listing instructions vs {len(data)} profiled', file=sys.stderr)
    tot = sum(int(r[ci]) for r in data); tots = sum(int(r[si]) for r in data)
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])
    for r, lc in zip(data, locs):
        a = agg[lc]
        a[0] += int(r[ci]); a[1] += int(r[si]); a[2] += int(r[ti]); a[3] += 1
    print(f'total warp inst {tot}  stall samples {tots}')
    print('inst%  samp%  thr/inst  n_sass  location')
    for lc, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if a[0] == 0 and a[1] == 0: continue
        print(f'{100*a[0]/tot:6.2f} {100*a[1]/max(1,tots):6.2f} {a[2]/max(1,a[0]):6.1f} {a[3]:5d}  {lc}')

main()
