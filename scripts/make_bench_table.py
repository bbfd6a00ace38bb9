"""Markdown tables from the committed bench lines (profiles/r2/bench_<config>_<N>gpu.json, one JSON line each, produced by
`bench.py --gpus N --steps 20 --warmup 5`): python scripts/make_bench_table.py > profiles/r2/bench_results.md"""
import glob
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = {"pa": "1. PA binary, 28-d, Asynchronous", "orr": "2. ORR 1 k-d, Synchronous (bench default)",
         "fgm": "3. RegressorPA + PolynomialFeatures(2) + StandardScaler, FGM", "kmeans": "4. K-means K = 256, 128-d, SSP",
         "mcpa": "5. MultiClassPA 100 k-d sparse (10 classes, 128 non-zeros), Synchronous"}


def load():
    out = {}
    for f in glob.glob(os.path.join(ROOT, "profiles", "r2", "bench_*gpu*.json")):
        base = os.path.basename(f)[len("bench_"):-len(".json")]
        cfg, n = base.split("_")[0], base.split("_")[1]
        try:
            j = json.loads(open(f).read().strip().splitlines()[-1])
        except Exception:
            continue
        out[(cfg, int(n.replace("gpu", "")))] = j
    return out


def main():
    R = load()
    print("# Round-2 bench lines (B200, `bench.py --gpus N --steps 20 --warmup 5`, CUDA events, max over ranks, median window)\n")
    print("Files: `profiles/r2/bench_<config>_<N>gpu.json` (the full JSON lines: windows, clocks, e2e block, baseline block, "
          "sync_check).  `vs baseline` = ours / the in-repo PyTorch + NCCL (+ cuBLAS) arm measured in the same invocation; it "
          "is not the reference (which cannot run here).\n")
    print("| config | GPUs | records/s | µs per step | vs baseline | baseline records/s | e2e records/s | e2e H2D GB/s per GPU "
          "(fraction of the box's pinned→device ceiling) | sync_check | SM MHz (reasons) |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for cfg in ("orr", "pa", "fgm", "kmeans", "mcpa"):
        for n in (1, 2, 4, 8):
            j = R.get((cfg, n))
            if not j:
                continue
            e = j.get("e2e") or {}
            b = j.get("baseline") or {}
            sc = j.get("sync_check") or {}
            clk = j.get("clocks") or {}
            frac = e.get("fraction_of_h2d_ceiling")
            print(f"| {NAMES[cfg]} | {n} | {j['value'] / 1e6:,.1f} M | {j['ms_per_step'] * 1e3:.2f} | "
                  f"{(j.get('vs_baseline') or 0):.1f}× | {(b.get('value') or 0) / 1e6:,.1f} M | "
                  f"{(e.get('value') or 0) / 1e6:,.1f} M | {(e.get('h2d_gbs_per_gpu') or 0):.1f}"
                  f"{f' ({frac:.2f})' if frac else ''} | {sc.get('ok')} | {clk.get('sm_mhz')} ({', '.join(clk.get('reasons') or []) or 'none'}) |")
    o1 = R.get(("orr", 1))
    if o1:
        print("\nWeak scaling of the flagship (per-GPU batch fixed at 32 768 records): efficiency = records/s(N) / (N × records/s(1))\n")
        print("| GPUs | records/s | µs per step | efficiency |")
        print("|---|---|---|---|")
        for n in (1, 2, 4, 8):
            j = R.get(("orr", n))
            if j:
                print(f"| {n} | {j['value'] / 1e6:,.1f} M | {j['ms_per_step'] * 1e3:.2f} | {j['value'] / (n * o1['value']):.3f} |")


if __name__ == "__main__":
    main()
