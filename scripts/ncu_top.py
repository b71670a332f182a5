"""Top stalled SASS instructions of one kernel in an ncu report: ncu_top.py REPORT KERNEL_REGEX [N]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + pat, "-c", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if 'Source' in r and '# Samples' in r)
h = rows[hi]
data = []
for r in rows[hi + 1:]:
    if len(r) != len(h) or r == h:
        break
    data.append(r)
ia = h.index('Source'); isamp = h.index('# Samples'); iex = h.index('Instructions Executed')
cols = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
tot = sum(int(r[isamp] or 0) for r in data)
print("samples", tot, "instructions", sum(int(r[iex] or 0) for r in data), "sass lines", len(data))
order = sorted(range(len(data)), key=lambda i: -int(data[i][isamp] or 0))[:n]
for i in sorted(order):
    r = data[i]
    st = {c[6:]: int(r[h.index(c)] or 0) for c in cols}
    st = {k: v for k, v in st.items() if v > tot * 0.004}
    print("%4d %5.1f%% %9s  %-60s %s" % (i, 100 * int(r[isamp] or 0) / max(tot, 1), r[iex], r[ia][:60], st))
