import csv, sys
keys = ["gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","dram__throughput.avg.pct_of_peak_sustained_elapsed",
"sm__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
"l1tex__data_pipe_lsu_wavefronts_mem_shared.sum","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
"sm__inst_executed.sum","smsp__inst_executed.sum","sm__warps_active.avg.pct_of_peak_sustained_active","launch__registers_per_thread",
"launch__occupancy_limit_registers","launch__occupancy_limit_shared_mem","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
"sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active","smsp__issue_active.avg.pct_of_peak_sustained_active","launch__grid_size","launch__block_size",
"launch__shared_mem_per_block_dynamic","sm__cycles_elapsed.avg","lts__t_sector_hit_rate.pct","smsp__cycles_active.avg"]
for f in sys.argv[1:]:
    rows = list(csv.reader(open(f)))
    hdr = rows[0]; units = rows[1]; val = rows[2]
    print("==", f, val[hdr.index("Kernel Name")][:60])
    for k in keys:
        for i,h in enumerate(hdr):
            if h == k: print(f"  {k:75s} {val[i]:>18s} {units[i]}")
    for i,h in enumerate(hdr):
        if "warp_issue_stalled" in h and h.endswith("_per_warp_active.pct") and "not_issued" not in h:
            try:
                if float(val[i].replace(",",""))>4: print(f"  {h:75s} {val[i]:>18s}")
            except: pass
