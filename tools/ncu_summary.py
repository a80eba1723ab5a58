"""Key raw-page metrics per captured launch of one or more .ncu-rep files.
    python tools/ncu_summary.py gpurun_out/a.ncu-rep [b.ncu-rep ...] > profiles/<tag>_ncu_full_summary.txt"""
import csv, io, subprocess, sys
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__cycles_active.avg", "launch__grid_size", "launch__block_size",
        "smsp__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_tensor.sum"]
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = {n: i for i, n in enumerate(hdr)}
    print("===", rep.split("/")[-1])
    for r in rows[2:]:
        print("----")
        print("  Kernel Name =", r[idx["Kernel Name"]][:110])
        for k in KEYS:
            if k in idx:
                print(f"  {k} [{units[idx[k]]}] = {r[idx[k]]}")
