"""Turn a gpurun_out/*.ncu-rep capture into the committed summaries under profiles/."""
import csv, io, subprocess, sys, os
rep = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/prof_orr.ncu-rep"
tag = sys.argv[2] if len(sys.argv) > 2 else "orr_step"
os.makedirs("profiles", exist_ok=True)
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keep = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "smsp__cycles_active.avg"]
idx = [hdr.index(k) for k in keep if k in hdr]
with open(f"profiles/{tag}_ncu_summary.csv", "w", newline="") as f:
    w = csv.writer(f)
    w.writerow([hdr[i] for i in idx]); w.writerow([units[i] for i in idx])
    for r in rows[2:]:
        w.writerow([r[i] for i in idx])
# per-kernel top stall lines from the source page
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
out, kernel, table, seen = [], None, [], set()
def flush():
    if kernel and table and kernel not in seen:
        seen.add(kernel)
        tot = sum(t[0] for t in table) or 1
        out.append(f"## {kernel}\n total stall samples {tot}\n")
        for smp, ins in sorted(table, reverse=True)[:14]:
            out.append(f"  {100.0 * smp / tot:5.1f}%  {ins.strip()}")
        out.append("")
for r in csv.reader(io.StringIO(src)):
    if not r: continue
    if r[0] == "Kernel Name":
        flush(); kernel, table = r[1][:110], []
    elif r[0].startswith("0x"):
        try: table.append((int(r[2]), r[1]))
        except ValueError: pass
flush()
open(f"profiles/{tag}_stall_hotspots.txt", "w").write("\n".join(out))
print(open(f"profiles/{tag}_ncu_summary.csv").read())
print("\n".join(out)[:3000])
