python scripts/prof_planner.py > gpurun_out/prof_planner_run.log 2>&1 || { tail -5 gpurun_out/prof_planner_run.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_planner.csv python scripts/prof_planner.py > gpurun_out/ncu_planner.log 2>&1
tail -3 gpurun_out/prof_planner_run.log
