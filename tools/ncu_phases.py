"""Sample share per barrier-delimited segment of a kernel from an ncu report's source page.
usage: python tools/ncu_phases.py report.ncu-rep"""
import collections, csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
hi = [i for i, x in enumerate(r) if 'Source' in x][0]
h = r[hi]; si = h.index('Source'); k = [i for i, n in enumerate(h) if 'Samples' in n][0]
rows = []
for x in r[hi + 1:]:
    try: rows.append((int(x[k]), x[si]))
    except (ValueError, IndexError): pass
tot = sum(a for a, _ in rows)
bars = [i for i, (a, s) in enumerate(rows) if 'BAR.SYNC' in s]
prev = 0
for b in bars + [len(rows) - 1]:
    seg = rows[prev:b + 1]
    c = collections.Counter()
    for a, s in seg:
        t = s.split()
        c[t[1] if t[0].startswith('@') else t[0]] += a
    print(prev, b, round(100 * sum(a for a, _ in seg) / tot, 1), [(n, round(100 * v / tot, 1)) for n, v in c.most_common(5)])
    prev = b + 1
