"""Brief of one `ncu --set full` capture: python tools/ncu_brief.py file.ncu-rep  (time, DRAM / L2 / L1 traffic, local memory,
occupancy, lanes per instruction, stall reasons per issue)."""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
for v in rows[2:]:
    d = {n: (v[i], u[i]) for i, n in enumerate(h)}
    print("==", d.get("Kernel Name", ("?",))[0][:100])
    keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__grid_size", "launch__occupancy_limit_registers",
            "launch__occupancy_limit_shared_mem", "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
            "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum",
            "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum", "smsp__sass_inst_executed_op_local_ld.sum",
            "smsp__sass_inst_executed_op_local_st.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
            "lts__t_sectors_srcunit_tex_op_write.sum", "l1tex__data_pipe_lsu_wavefronts.sum",
            "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
            "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed", "l1tex__f_wavefronts.sum",
            "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum", "l1tex__data_bank_reads.avg.pct_of_peak_sustained_elapsed",
            "l1tex__lsuin_requests.avg.pct_of_peak_sustained_elapsed", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]
    for k in keys:
        if k in d:
            print(f"{k:80s} {d[k][0]} {d[k][1]}")
    for n in d:
        if "issue_stalled" in n and n.endswith("per_issue_active.ratio") and float(d[n][0] or 0) >= 0.3:
            print(f"{n:80s} {d[n][0]}")
