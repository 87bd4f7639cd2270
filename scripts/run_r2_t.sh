#!/bin/bash
# round 2, call T (group kNN kernel, lane-per-sphere state checks, allocation-free scan entry point): GPU tests, smoke, bench (both arms), ncu launch list + full captures of the three hot kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2t_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2t_pytest.log; tail -5 gpurun_out/r2t_pytest.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/r2t_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2t_smoke.log
timeout 900 python bench.py > gpurun_out/r2t_bench.json 2> gpurun_out/r2t_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2t_bench.err
timeout 600 python bench.py --impl reference --steps 5 > gpurun_out/r2t_bench_ref.json 2>> gpurun_out/r2t_bench.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2t_launches.csv python bench.py --steps 3 --warmup 3 --skip-cpu --skip-planner --skip-sharded > gpurun_out/r2t_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:grid_knn -o gpurun_out/r2t_prof_knn -f python scripts/prof_run.py knn_grid > gpurun_out/r2t_ncu_knn.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:edges_|collide_edges|collide_spheres" --launch-skip 4 -o gpurun_out/r2t_prof_edges -f python scripts/prof_run.py edges spheres > gpurun_out/r2t_ncu_edges.log 2>&1
ls -la gpurun_out | tail -20
