# full single-GPU check: GPU tests, smoke, bench (both arms), ncu launch list + full captures
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench_k.json 2> gpurun_out/bench_k.err; echo "bench rc=$?" >> gpurun_out/bench_k.err
python bench.py --impl reference --steps 3 > gpurun_out/bench_ref_k.json 2>> gpurun_out/bench_k.err
python scripts/ab_e2e.py > gpurun_out/ab_e2e_k.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_k.csv python bench.py --steps 3 --warmup 3 --skip-cpu --skip-planner > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:grid_knn -o gpurun_out/prof_knn_k -f python scripts/prof_run.py knn_grid > gpurun_out/ncu_knn_k.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:edges_|collide_edges|collide_spheres" --launch-skip 4 -o gpurun_out/prof_edges_k -f python scripts/prof_run.py edges spheres > gpurun_out/ncu_edges_k.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; tail -2 gpurun_out/bench_k.err; tail -1 gpurun_out/ab_e2e_k.log
