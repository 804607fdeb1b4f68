"""Pull the counters the roofline discussion needs out of an ncu report (run here, no GPU needed):
    python scripts/ncu_summary.py gpurun_out/r3_search.ncu-rep [more.ncu-rep ...]
Prints one block per report: duration, DRAM bytes read / written, DRAM throughput, issue / warp activity, L2 hit rate,
registers, shared memory, grid — the numbers profiles/*_ncu_summary.txt and *_traffic.json quote."""
import csv
import io
import json
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__grid_size",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_bytes.sum", "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]

for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        print(rep, "no data")
        continue
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        rec = dict(zip(hdr, r))
        print(f"# {rep}: {rec.get('Kernel Name', '?')}")
        got = {}
        for k in WANT:
            if k in rec:
                u = units[hdr.index(k)]
                print(f"{k:72s} {u:18s} {rec[k]}")
                got[k] = (rec[k], u)
        print(json.dumps({k: v[0] + " " + v[1] for k, v in got.items()}))
