"""Print the last step's kernels (name, ms) from an ncu gpu__time_duration CSV."""
import csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value')
seq = [(re.sub(r'\(.*', '', r[ki]).replace('mtg::<unnamed>::', '').replace('void ', ''), float(r[vi].replace(',', ''))) for r in rows[1:]]
idx = [i for i, s in enumerate(seq) if "extract_kernel" in s[0] or "extract_suffix" in s[0]]
start = idx[-1] if idx else 0
tot = 0
prev = None; acc = 0; cnt = 0
def flush():
    if prev is not None: print("%-70s x%-3d %8.3f ms" % (prev[:70], cnt, acc / 1e6))
for name, ns in seq[start:]:
    tot += ns
    if name == prev: acc += ns; cnt += 1
    else:
        flush(); prev = name; acc = ns; cnt = 1
flush()
print("total %.3f ms" % (tot / 1e6))
