"""Summarise an `ncu -i <rep> --page raw --csv` dump: one row per launch (duration, DRAM traffic, tensor pipe, L2 hit rate),
optionally restricted to ONE iteration (the launches from `--first` up to the next launch of the same kernel), plus totals.

    python tools/ncu_summary.py raw.csv "title" [--first filter_pack_batch_kernel] [--labels a,b,c] [--traffic-json out.json key source]
"""
import csv
import json
import sys

M = {
    "dur": "gpu__time_duration.sum",
    "rd": "dram__bytes_read.sum",
    "wr": "dram__bytes_write.sum",
    "tc": "sm__ops_path_tensor_op_utchmma_src_tf32_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "l2": "lts__t_sector_hit_rate.pct",
    "grid": "launch__grid_size",
    "regs": "launch__registers_per_thread",
    "sm": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "occ": "sm__warps_active.avg.pct_of_peak_sustained_active",
}
SCALE = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def main():
    args = sys.argv[1:]
    path, title = args[0], args[1]
    first = args[args.index("--first") + 1] if "--first" in args else None
    labels = args[args.index("--labels") + 1].split(",") if "--labels" in args else []
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {k: hdr.index(v) for k, v in M.items() if v in hdr}
    ki = hdr.index("Kernel Name")

    def val(r, k):
        if k not in col:
            return float("nan")
        s = r[col[k]].replace(",", "")
        try:
            return float(s) * SCALE.get(units[col[k]], 1.0)
        except ValueError:
            return float("nan")
    names = [r[ki].split("(")[0].replace("void ", "").replace("sn::", "") for r in data]
    lo, hi = 0, len(data)
    if first:
        idx = [i for i, n in enumerate(names) if n.startswith(first)]
        if len(idx) >= 2:
            lo, hi = idx[0], idx[1]
        elif idx:
            lo = idx[0]
    print(f"# {title}\n")
    print("| # | kernel | what | grid | duration us | DRAM read MB | DRAM write MB | (read+write)/duration GB/s | tensor pipe % (elapsed) | "
          "SM busy % | occupancy % | L2 hit % | regs |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    tot = {"dur": 0.0, "rd": 0.0, "wr": 0.0}
    for n, i in enumerate(range(lo, hi)):
        r = data[i]
        d, rd, wr = val(r, "dur"), val(r, "rd"), val(r, "wr")
        tot["dur"] += d
        tot["rd"] += rd
        tot["wr"] += wr
        lab = labels[n] if n < len(labels) else ""
        print(f"| {n} | `{names[i]}` | {lab} | {int(val(r, 'grid'))} | {d:.1f} | {rd:.1f} | {wr:.1f} | {(rd + wr) / d * 1e3:.0f} | "
              f"{val(r, 'tc'):.1f} | {val(r, 'sm'):.0f} | {val(r, 'occ'):.0f} | {val(r, 'l2'):.1f} | {int(val(r, 'regs'))} |")
    print(f"\nTotals over these {hi - lo} launches: {tot['dur']:.0f} us (serialised, cold caches), DRAM read {tot['rd']:.1f} MB + write "
          f"{tot['wr']:.1f} MB = **{tot['rd'] + tot['wr']:.1f} MB**.")
    if "--traffic-json" in args:
        k = args.index("--traffic-json")
        out, key, source = args[k + 1], args[k + 2], args[k + 3]
        try:
            t = json.load(open(out))
        except (OSError, ValueError):
            t = {}
        t[key] = {"bytes": (tot["rd"] + tot["wr"]) * 1e6, "launches": hi - lo, "source": source}
        json.dump(t, open(out, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
