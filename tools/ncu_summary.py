"""Condenses an ncu report (`ncu --set full`) into the text summary kept under profiles/: one block per captured launch
with the metrics the design discussion uses (duration, DRAM bytes, pipe utilisation, occupancy limits, stall reasons).
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/ncu_<kernel>_<round>_summary.txt"""
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(r"^(Kernel Name|dram__bytes_(read|write)\.sum($|\.per_second)|gpu__dram_throughput|gpu__time_duration\.sum|"
                  r"l1tex__data_pipe_tc_wavefronts_mem_shared\.sum\.pct|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$|"
                  r"launch__(block_size|cluster_size|grid_size|occupancy_limit_\w+|registers_per_thread|shared_mem_per_block_dynamic|"
                  r"waves_per_multiprocessor)$|lts__t_sector_hit_rate\.pct|lts__throughput\.avg\.pct|sm__cycles_active\.avg$|"
                  r"sm__inst_executed\.avg\.per_cycle_active|sm__inst_executed_pipe_(alu|fma|xu|uniform)\.avg\.pct_of_peak_sustained_active|"
                  r"sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)|sm__throughput\.avg\.pct|"
                  r"sm__warps_active\.avg\.pct|smsp__issue_active\.avg\.pct|smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio|"
                  r"smsp__cycles_active\.avg$|sm__cycles_elapsed\.avg\.per_second)")


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for n, r in enumerate(rows[2:]):
        print(f"=== launch {n} of {path.split('/')[-1]}")
        for h, u, v in zip(hdr, units, r):
            if KEEP.match(h) and v != "":
                print(f"{h} [{u}] = {v}")


if __name__ == "__main__":
    main(sys.argv[1])
