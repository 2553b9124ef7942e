"""Summarise an .ncu-rep (run where ncu is installed, no GPU needed): key raw metrics per captured launch.
usage: python tools/ncu_read.py gpurun_out/x.ncu-rep [more.ncu-rep ...]"""
import csv, io, subprocess, sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__inst_executed_pipe_xu.sum", "sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.sum", "sm__cycles_elapsed.avg", "sm__cycles_active.avg",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio" ]

def main():
    for path in sys.argv[1:]:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hdr = rows[0]
        print("==", path)
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            print("--", d.get("Kernel Name", "")[:90], "grid", d.get("Grid Size"), "block", d.get("Block Size"))
            for k in hdr:
                kk = k.lower()
                if any(s in kk for s in ("gpu__time_duration.sum", "registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                         "dram__throughput.avg.pct", "sm__throughput.avg.pct", "pipe_tensor", "pipe_xu", "issue_active.avg.pct",
                                         "warps_active.avg.pct", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct", "lts__throughput.avg.pct",
                                         "occupancy_limit", "sm__warps_active", "achieved_occupancy", "pipe_fma_cycles_active.avg.pct", "pipe_alu_cycles_active.avg.pct",
                                         "mem_shared.sum", "bank_conflicts", "smsp__inst_executed.sum", "launch__waves")):
                    if d[k] not in ("", "n/a"):
                        print(f"   {k:95s} {d[k]}")

if __name__ == "__main__":
    main()
