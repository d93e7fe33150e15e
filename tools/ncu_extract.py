"""Condense an .ncu-rep (ncu --set full) into a text table: one block of key metrics per kernel instance.
    python tools/ncu_extract.py gpurun_out/prof.ncu-rep > profiles/ncu_xxx.txt"""
import csv
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__registers_per_thread", "regs/thread"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem/block"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor instr"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
    ("lts__t_bytes.sum", "L2 bytes"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1/TEX throughput %"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    ("smsp__cycles_active.avg", "SM active cycles"), ("sm__cycles_elapsed.max", "elapsed cycles"),
]
print(f"# extracted from {rep} (ncu --set full --clock-control none); one block per captured launch")
for r in data:
    if len(r) < len(hdr):
        continue
    name = r[ix["Kernel Name"]].replace("tadnn::<unnamed>::", "").replace("unnamed>::", "")
    print(f"\n== {name[:110]}")
    for key, label in WANT:
        if key in ix:
            print(f"   {label:26s} {r[ix[key]]:>16s} {units[ix[key]]}")
