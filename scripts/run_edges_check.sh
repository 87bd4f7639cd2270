python -m pytest tests/test_gpu_collide.py tests/test_gpu_planner.py -m gpu -x -q > gpurun_out/pytest_edges.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_edges.log
python bench.py --steps 20 --skip-cpu > gpurun_out/bench_d.json 2> gpurun_out/bench_d.err
ncu --set full --clock-control none --import-source on -k regex:edges_ -o gpurun_out/prof_edges_d -f python scripts/prof_run.py edges > gpurun_out/ncu_edges_d.log 2>&1
tail -3 gpurun_out/pytest_edges.log
