"""Print selected metrics per kernel from an `ncu --page raw --csv` dump.  usage: ncu_pick.py raw.csv [substr ...]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
DEFAULT = ["gpu__time_duration.sum", "dram__bytes_read.sum [", "dram__bytes_write.sum [", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__throughput.avg.pct", "lts__throughput.avg.pct", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct",
           "smsp__issue_active.avg.pct", "sm__inst_executed_pipe_tmem.avg.pct", "sm__pipe_tensor_cycles_active_realtime.avg.pct",
           "pipe_tensor_subpipe_hmma_cycles_active_realtime", "launch__registers_per_thread [", "smsp__inst_executed.sum [",
           "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum [", "sm__warps_active.avg.pct", "smsp__cycles_active.avg [",
           "sm__cycles_elapsed.avg [", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum [", "smsp__warp_issue_stalled", "lts__t_bytes.sum [",
           "l1tex__t_bytes.sum [", "sm__inst_executed_pipe_lsu", "sm__inst_executed_pipe_alu.avg.pct", "sm__inst_executed_pipe_fma.avg.pct",
           "sm__inst_executed_pipe_xu.avg.pct", "smsp__average_warp", "smsp__pcsamp_warps_issue_stalled"]
pats = sys.argv[2:] or DEFAULT
ki = hdr.index("Kernel Name")
for r in rows[2:]:
    print("=====", r[ki][:90])
    for i, h in enumerate(hdr):
        key = f"{h} [{units[i]}"
        if any(p in key + "]" or p in key for p in pats) and r[i] not in ("", "0", "0.000000"):
            print(f"  {h} [{units[i]}] = {r[i]}")
