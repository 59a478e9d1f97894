#!/usr/bin/env python
"""Attribute executed SASS instructions of one kernel to source lines.
  tools/ncu_by_line.py <ncu source-page csv> <nvdisasm --print-line-info txt> <mangled kernel name> [launch idx]
The ncu source page lists SASS in address order, nvdisasm lists the same instructions with
//## File ... line N markers; the two are zipped and checked opcode by opcode."""
import csv
import re
import sys
from collections import Counter


def main():
    src_csv, dis_txt, mangled = sys.argv[1:4]
    launch = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    rows = list(csv.reader(open(src_csv)))
    starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name'] + [len(rows)]
    s, e = starts[launch], starts[launch + 1]
    hdr = rows[s + 1]
    isrc, iex, ismp = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
    body = rows[s + 2:e]
    # nvdisasm section of the kernel
    lines = open(dis_txt).read().split('\n')
    begin = next(i for i, l in enumerate(lines) if l.startswith('.text.' + mangled + ':'))
    cur, ins = None, []
    for l in lines[begin + 1:]:
        if l.startswith('.text.') or l.startswith('.section'):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            if 'inlined at' not in l:
                cur = (m.group(1).split('/')[-1], int(m.group(2)))
            continue
        m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*?);', l)
        if m:
            ins.append((cur, m.group(2)))
    assert len(ins) == len(body), (len(ins), len(body))
    per_line, per_line_s = Counter(), Counter()
    tot = ts = 0
    for (loc, text), r in zip(ins, body):
        n, sm = int(r[iex]), int(r[ismp])
        per_line[loc] += n
        per_line_s[loc] += sm
        tot += n
        ts += sm
    print(f"total executed {tot}, samples {ts}")
    by_samples = len(sys.argv) > 5 and sys.argv[5] == 'samples'
    order = per_line_s.most_common(40) if by_samples else per_line.most_common(60)
    for loc, _ in order:
        n = per_line[loc]
        print(f"{str(loc):40s} exec {n:11d} {100*n/tot:5.1f}%   samples {100*per_line_s[loc]/max(ts,1):5.1f}%")


if __name__ == '__main__':
    main()
