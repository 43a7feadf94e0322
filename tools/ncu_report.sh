#!/bin/bash
# usage: tools/ncu_report.sh <name> <function-pattern> <phase-ranges> [part]   (reads gpurun_out/<name>.ncu-rep;
# part = the avr_volpath_p<part>.o object that holds the kernel: 2*ZSOBOL + EXACT)
set -e
cd "$(dirname "$0")/.."
n=$1; part=${4:-0}
ncu -i gpurun_out/$n.ncu-rep --page raw --csv > gpurun_out/$n.raw.csv 2>/dev/null
ncu -i gpurun_out/$n.ncu-rep --page source --csv > gpurun_out/$n.src.csv 2>/dev/null
mkdir -p gpurun_out/cubin; (cd gpurun_out/cubin && rm -f *.cubin && cuobjdump -xelf all ../../acceleratedvolrenderer_b200/csrc/avr_volpath_p$part.o >/dev/null && nvdisasm -gi -c *.sm_100a.cubin > disgi.txt 2>&1)
python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/$n.raw.csv')))
hdr=rows[0]; vals=rows[2]
want=['gpu__time_duration.sum','smsp__thread_inst_executed_per_inst_executed.ratio','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct','smsp__average_warps_issue_stalled','launch__grid_size','smsp__inst_executed.sum','dram__bytes_read.sum','dram__bytes_write.sum','launch__shared_mem_per_block_dynamic','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','sm__cycles_elapsed.max','smsp__warps_eligible.avg.per_cycle_active','lts__t_bytes.sum ','l1tex__t_bytes.sum ']
for h,v in zip(hdr,vals):
    if any(h.startswith(w) for w in want) and 'per_second' not in h:
        try:
            if float(v.replace(',',''))==0: continue
        except: pass
        print(h,v)
PY
python tools/ncu_by_line.py gpurun_out/$n.src.csv gpurun_out/cubin/disgi.txt "$2" 0 "$3" 2>/dev/null | sed -n '/by phase/,$p'
