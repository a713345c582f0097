#!/bin/bash
mkdir -p gpurun_out/r02m
python tools/limit_probe.py > gpurun_out/r02m/limit_probe.txt 2>&1; cat gpurun_out/r02m/limit_probe.txt
