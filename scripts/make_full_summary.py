"""Summarise the `ncu --set full` capture of scripts/prof_full.sh: one row per kernel launch of one timed window of
each bench configuration -> profiles/r2/full_set_rows.csv (selected raw metrics) and full_set_summary.md (one row per
kernel and configuration with the roofline fractions against MEASURED_PEAKS.json).
usage: python scripts/make_full_summary.py gpurun_out/prof/full_all_raw.csv
       python scripts/make_full_summary.py gpurun_out/prof/every_kernel_raw.csv every_kernel
           (capture of scripts/prof_every_kernel.sh: one row per kernel function at the tests' shapes, no configs)"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor"]
SPECIFIC = [("orr", ("gram_f16", "sq_residual", "ridge_", "mirror_upper", "sync_gram")),
            ("pa", ("linear_minibatch", "linear_exact")),
            ("fgm", ("poly_", "fused_commit")),
            ("kmeans", ("umma_kmeans", "kmeans_", "umma_gemm")),
            ("mcpa", ("mcpa_", "sparse_rows"))]
STARTERS = ("holdout_swap", "sparse_rows", "expand_rows", "stage_rows")


def short(name):
    return name.split("(")[0].replace("omldm::", "").replace("void ", "")[:60]


def config_of(name):
    for cfg, keys in SPECIFIC:
        if any(k in name for k in keys):
            return cfg
    return None


def to_bytes(v, unit):
    f = float(v.replace(",", "")) if v else 0.0
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit.split("/")[0], 1)


def to_us(v, unit):
    f = float(v.replace(",", "")) if v else 0.0
    return f * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(unit, 1)


def main():
    path = sys.argv[1]
    stem = sys.argv[2] if len(sys.argv) > 2 else "full_set"
    per_config = stem == "full_set"
    rows = list(csv.reader(open(path)))
    H, U, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(H)}
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    hbm = float(peaks["hbm_gbs"])
    names = [short(r[ix["Kernel Name"]]) for r in data]
    cfgs = [config_of(n) if per_config else "-" for n in names]
    for i in range(len(data) if per_config else 0):  # rows without a configuration-specific name
        if cfgs[i] is None:
            order = list(range(i + 1, len(data))) if any(s in names[i] for s in STARTERS) else list(range(i - 1, -1, -1))
            for j in order:
                if config_of(names[j]):
                    cfgs[i] = config_of(names[j])
                    break
    out_csv = os.path.join(ROOT, "profiles", "r2", stem + "_rows.csv")
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "config", "kernel"] + [c + " [" + U[ix[c]] + "]" for c in COLS if c in ix])
        for r, n, c in zip(data, names, cfgs):
            w.writerow([r[ix["ID"]], c, n] + [r[ix[k]] for k in COLS if k in ix])
    agg = {}
    for r, n, c in zip(data, names, cfgs):
        agg.setdefault((c, n), []).append(r)
    lines = ["| config | kernel | launches in the window | time µs | grid x block | regs | smem KB (dyn + static) | DRAM MB (r + w) | "
             "DRAM GB/s | of measured HBM copy peak | tensor pipe % | SM busy % |", "|---|---|---|---|---|---|---|---|---|---|---|---|"]
    for (c, n), rs in agg.items():
        r = rs[-1]
        g = lambda k: r[ix[k]] if k in ix else ""
        us = to_us(g("gpu__time_duration.sum"), U[ix["gpu__time_duration.sum"]])
        rd = to_bytes(g("dram__bytes_read.sum"), U[ix["dram__bytes_read.sum"]])
        wr = to_bytes(g("dram__bytes_write.sum"), U[ix["dram__bytes_write.sum"]])
        gbs = (rd + wr) / us / 1e3 if us else 0.0
        dyn = to_bytes(g("launch__shared_mem_per_block_dynamic"), U[ix["launch__shared_mem_per_block_dynamic"]]) / 1e3
        sta = to_bytes(g("launch__shared_mem_per_block_static"), U[ix["launch__shared_mem_per_block_static"]]) / 1e3
        lines.append(f"| {c} | `{n}` | {len(rs)} | {us:.1f} | {g('launch__grid_size')} x {g('launch__block_size')} | "
                     f"{g('launch__registers_per_thread')} | {dyn:.1f} + {sta:.1f} | {rd / 1e6:.1f} + {wr / 1e6:.1f} | {gbs:.0f} | "
                     f"{gbs / hbm:.2f} | {float(g('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active') or 0):.1f} | "
                     f"{float(g('sm__throughput.avg.pct_of_peak_sustained_elapsed') or 0):.1f} |")
    out_md = os.path.join(ROOT, "profiles", "r2", stem + "_summary.md")
    with open(out_md, "w") as f:
        if not per_config:
            f.write("# `ncu --set full` row for every kernel function of the library (tests' shapes), 1 B200\n\n"
                    "Capture: `scripts/prof_every_kernel.sh` - the GPU tests under `ncu --set full --clock-control none "
                    "--kernel-name-base demangled --kernel-id ::regex:omldm:2` (second invocation of every kernel whose name "
                    "contains `omldm::`; protocol kernels run with emulated ranks inside one cooperative launch).  The shapes are "
                    "small: read registers, shared memory, grid and pipe mix here, and the bench-shape times in "
                    "`full_set_summary.md`.\n\n" + "\n".join(sorted(lines[2:], key=lambda l: l.split('`')[1]) if False else lines) + "\n")
            print("\n".join(lines))
            return
        f.write("# `ncu --set full` rows, one timed window (2 steps) of every bench configuration, 1 B200\n\n"
                "Capture: `scripts/prof_full.sh` (plain run first, then `ncu --set full --clock-control none "
                "--profile-from-start off` around the window; kernels run serialised with cold caches under ncu, so "
                "times are upper bounds of the in-step times and are never bench values).  Peak: "
                f"`MEASURED_PEAKS.json` hbm_gbs = {hbm:.1f}.  Last launch of each kernel; all launches with more raw "
                "metrics in `full_set_rows.csv`.\n\n" + "\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
