"""Key metrics of an .ncu-rep as a small CSV (the reports themselves are MBs and stay in gpurun_out/).
   python bench_tools/ncu_summary.py gpurun_out/prof_x.ncu-rep > profiles/ncu/prof_x_summary.csv"""
import csv
import subprocess
import sys

KEEP = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput", "gpu__dram_throughput",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate", "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__pipe_tensor_cycles_active",
        "sm__inst_executed_pipe_tensor", "sm__warps_active", "sm__throughput", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__cluster", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__warp_issue_stalled")
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [f"launch {i}" for i in range(len(rows) - 2)])
for j, h in enumerate(hdr):
    if h.startswith(KEEP):
        w.writerow([h, units[j]] + [r[j] for r in rows[2:]])
