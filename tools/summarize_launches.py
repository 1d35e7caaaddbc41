"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel."""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
hdr = rows[hi]; data = rows[hi + 1:]
ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in data:
    if len(r) <= vi: continue
    name = r[ki]; v = float(r[vi].replace(',', ''))
    if r[ui] == 'ns': v /= 1e3
    elif r[ui] == 'ms': v *= 1e3
    short = re.sub(r'\(.*', '', name)
    m = re.search(r'gemm_tc_kernel<(.+?), *(?:sn::)?(TEpi\w+), *\(int\)(\d+), *\(bool\)(\d)', name)
    if m: short = 'gemm_tc_kernel<' + m.group(1).replace('sn::', '').replace('(bool)', '') + ', ' + m.group(2) + ', BN=' + m.group(3) + (', split' if m.group(4) == '1' else '') + '>'
    agg[short][0] += 1; agg[short][1] += v
tot = sum(v[1] for v in agg.values())
title = sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]
print("# %s\n" % title)
print("Command (after the same command exited 0 without ncu): `ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline`  ")
print("Per-launch times are cold-cache and serialised: compare SHARES, not absolutes.  %d launches, %.1f ms total.\n" % (sum(v[0] for v in agg.values()), tot / 1e3))
print("| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print("| `%s` | %d | %.0f | %.1f%% | %.1f |" % (k[:110], v[0], v[1], 100 * v[1] / tot, v[1] / v[0]))
