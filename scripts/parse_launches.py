"""Summarise an ncu launch list (scripts/prof_launches.sh): per kernel name the launch count, the mean device time and the
share of the total, plus DRAM bytes, grid, registers and tensor-pipe activity of the last launch.
usage: python scripts/parse_launches.py gpurun_out/prof/orr.csv [skip_first_n]"""
import collections
import csv
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    H = rows[hdr]
    ik, im, iv, iid = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("ID")
    d = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) > iv:
            d.setdefault(r[iid], {"k": r[ik]})[r[im]] = r[iv]
    return [v for _, v in d.items()]


def num(v, key):
    try:
        return float(str(v.get(key, "0")).replace(",", ""))
    except ValueError:
        return 0.0


def main():
    seq = load(sys.argv[1])
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else len(seq) // 2
    seq = seq[skip:]
    agg = collections.OrderedDict()
    for v in seq:
        name = v["k"].split("(")[0].replace("omldm::", "")[:48]
        a = agg.setdefault(name, {"n": 0, "t": 0.0, "last": v})
        a["n"] += 1
        a["t"] += num(v, "gpu__time_duration.sum") / 1000.0
        a["last"] = v
    total = sum(a["t"] for a in agg.values())
    print(f"{'kernel':50s} {'n':>4s} {'us/launch':>10s} {'share':>6s} {'dram MB':>8s} {'grid':>6s} {'regs':>5s} {'tensor%':>8s}")
    for name, a in agg.items():
        l = a["last"]
        mb = (num(l, "dram__bytes_read.sum") + num(l, "dram__bytes_write.sum")) / 1e6
        print(f"{name:50s} {a['n']:4d} {a['t'] / a['n']:10.1f} {100 * a['t'] / total:5.1f}% {mb:8.1f} "
              f"{int(num(l, 'launch__grid_size')):6d} {int(num(l, 'launch__registers_per_thread')):5d} "
              f"{num(l, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'):8.1f}")


if __name__ == "__main__":
    main()
