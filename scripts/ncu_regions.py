"""Summarise an ncu report's source page: stall samples / instructions per SASS region."""
import csv, subprocess, sys, collections
rep = sys.argv[1]
step = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[1]; data = rows[2:]
ia = h.index('Source'); isamp = h.index('# Samples'); iex = h.index('Instructions Executed')
cols = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
idx = {c: h.index(c) for c in cols}
tot = sum(int(r[isamp] or 0) for r in data)
print("total samples", tot, "instr", sum(int(r[iex] or 0) for r in data), "lines", len(data))
for a in range(0, len(data), step):
    seg = data[a:a + step]
    s = sum(int(r[isamp] or 0) for r in seg); e = sum(int(r[iex] or 0) for r in seg)
    acc = collections.Counter()
    for r in seg:
        for c in cols:
            acc[c[6:]] += int(r[idx[c]] or 0)
    top = max(seg, key=lambda r: int(r[isamp] or 0))
    print("%4d %5.1f%% %6dM %-60s | %s" % (a, 100 * s / max(tot, 1), e // 1000000,
          dict(acc.most_common(3)), top[ia][:45]))
