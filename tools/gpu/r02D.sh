#!/bin/bash
mkdir -p gpurun_out/r02D
python tools/ncu_target_dephased.py full > gpurun_out/r02D/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"exo_motor_kernel|exo_step_kernel" -s 700 -c 6 --csv --log-file gpurun_out/r02D/motor_steady.csv python tools/ncu_target_dephased.py full > gpurun_out/r02D/ncu.log 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02D/motor_steady.csv')))
i=[k for k,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i]
for r in rows[i+1:]:
    if len(r)<len(h): continue
    d=dict(zip(h,r))
    print(d['Kernel Name'][:40], d['Metric Name'], d['Metric Value'])
PY
