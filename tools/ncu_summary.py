import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[0]; units=rows[1]; data=rows[2:]
want=['Kernel Name','Grid Size','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','dram__throughput.avg.pct_of_peak_sustained_elapsed','lts__t_bytes.sum','lts__throughput.avg.pct_of_peak_sustained_elapsed','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_tensor.sum','sm__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','lts__t_sector_hit_rate.pct','sm__cycles_elapsed.max','lts__t_sectors_srcunit_tex_op_read.sum','lts__t_sectors_srcunit_tex_op_write.sum','smsp__cycles_active.avg']
want += sys.argv[2:]
for w in want:
    idx=[i for i,h in enumerate(hdr) if h==w]
    if not idx:
        idx=[i for i,h in enumerate(hdr) if w in h][:4]
    for i in idx:
        print(f"{hdr[i]:78s} {units[i]:10s}", [r[i][:34] for r in data])
