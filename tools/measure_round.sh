set -x
mkdir -p gpurun_out/r1b
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 400 python bench.py > gpurun_out/r1b/bench_v5.json 2> gpurun_out/r1b/bench_v5.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1b/bench_ref_v5.json 2> gpurun_out/r1b/bench_ref_v5.err
timeout 300 python tools/sweep_targets.py > gpurun_out/r1b/sweep_v7.json 2> gpurun_out/r1b/sweep_v7.txt
SWEEP_ALIAS=1 timeout 300 python tools/sweep_targets.py > gpurun_out/r1b/sweep_v7_alias.json 2> gpurun_out/r1b/sweep_v7_alias.txt
timeout 300 python tools/bench_envs.py > gpurun_out/r1b/bench_envs.json 2> gpurun_out/r1b/bench_envs.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1b/launches.csv python bench.py --steps 5 --warmup 3 --no-e2e > gpurun_out/r1b/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_rollout|k_select_windows|k_make_batch|k_target_rows" -s 22 -c 5 -o gpurun_out/r1b/step_full python bench.py --steps 3 --warmup 3 --no-e2e > gpurun_out/r1b/ncu_full.log 2>&1
tail -2 gpurun_out/r1b/ncu_full.log
cat gpurun_out/r1b/bench_v5.json | head -c 600
