# ncu full capture (with source) of the sweep kernel on the configs[2] workload
python scripts/prof_run.py edges > gpurun_out/prof_sweep_run.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:edges_sweep --launch-skip 1 --launch-count 1 -o gpurun_out/prof_sweep_f -f python scripts/prof_run.py edges > gpurun_out/ncu_sweep_f.log 2>&1
tail -3 gpurun_out/prof_sweep_run.log; tail -3 gpurun_out/ncu_sweep_f.log
