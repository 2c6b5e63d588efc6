"""Brief per-kernel digest of an ncu report (read on the CPU box): time, DRAM bytes, LSU wavefronts, issue, stalls.
    python tools/ncu_brief.py gpurun_out/x.ncu-rep [keys]"""
import csv, subprocess, sys
rep = sys.argv[1]
keys = float(sys.argv[2]) if len(sys.argv) > 2 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
W = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
     "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
     "l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.sum",
     "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum",
     "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
     "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum", "sm__inst_issued.avg.pct_of_peak_sustained_active",
     "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
     "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
     "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__cycles_active.sum",
     "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
     "l1tex__m_l1tex2xbar_write_bytes.sum", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("==", d.get("Kernel Name", "?")[:90])
    for w in W:
        if w in d:
            v = d[w]
            extra = ""
            try:
                f = float(v.replace(",", ""))
                if keys and (w.endswith(".sum") and "time" not in w and "bytes" not in w):
                    extra = f"   per key {f / keys:.3f}" + (f" (x32 = {32 * f / keys:.1f} thread-instr)" if "inst_executed" in w else "")
            except ValueError:
                pass
            print(f"  {w:75s} {v} {units[hdr.index(w)]}{extra}")
    for i, h in enumerate(hdr):
        if "issue_stalled" in h and h.endswith("per_issue_active.ratio"):
            try:
                f = float(r[i])
            except ValueError:
                continue
            if f > 0.3:
                print(f"  stall {h.split('issue_stalled_')[1].split('_per_issue')[0]:30s} {f:.2f}")
