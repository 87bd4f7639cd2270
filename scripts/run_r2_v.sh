#!/bin/bash
# round 2: launch lists of a host-driven planner step at the configs[3] shape (150 k nodes) and at the configs[4](i) shape (batch 1024, 8 k nodes)
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2v_launches_planner_cfg3.csv python scripts/prof_planner.py > gpurun_out/r2v_cfg3.log 2>&1; tail -2 gpurun_out/r2v_cfg3.log
BATCH=1024 GROW=8000 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2v_launches_planner_b1024.csv python scripts/prof_planner.py > gpurun_out/r2v_b1024.log 2>&1; tail -2 gpurun_out/r2v_b1024.log
