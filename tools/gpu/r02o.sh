#!/bin/bash
mkdir -p gpurun_out/r02o
python tools/ncu_target.py 65536 6 full > gpurun_out/r02o/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,launch__registers_per_thread,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none --csv --log-file gpurun_out/r02o/launches_motor.csv python tools/ncu_target.py 65536 6 full > gpurun_out/r02o/ncu.log 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02o/launches_motor.csv')))
i=[k for k,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i]
for r in rows[i+1:]:
    if len(r)<len(h): continue
    d=dict(zip(h,r))
    print(d['Kernel Name'][:60], d['Metric Name'], d['Metric Value'])
PY
