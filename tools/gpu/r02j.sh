#!/bin/bash
mkdir -p gpurun_out/r02j
python tools/step_timeline.py > gpurun_out/r02j/timeline_65536.txt 2>&1; cat gpurun_out/r02j/timeline_65536.txt
EXO_STEP_STAGGER=1500 python tools/step_timeline.py > gpurun_out/r02j/timeline_65536_stagger1500.txt 2>&1; cat gpurun_out/r02j/timeline_65536_stagger1500.txt
python tools/step_timeline.py 524288 > gpurun_out/r02j/timeline_524288.txt 2>&1; cat gpurun_out/r02j/timeline_524288.txt
