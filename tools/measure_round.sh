# One GPU box, final state of a session: tests, bench (both arms), sweeps, env rates, ncu launch list and --set full capture.
#   gpurun --timeout 1500 -- 'bash tools/measure_round.sh r1c'
set -x
TAG=${1:-r1c}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err
timeout 400 python bench.py --impl reference > $OUT/bench_ref.json 2> $OUT/bench_ref.err
SWEEP_CPU=1 timeout 600 python tools/sweep_targets.py > $OUT/sweep.json 2> $OUT/sweep.txt
SWEEP_ALIAS=1 timeout 300 python tools/sweep_targets.py > $OUT/sweep_alias.json 2> $OUT/sweep_alias.txt
timeout 300 python tools/bench_envs.py > $OUT/bench_envs.json 2> $OUT/bench_envs.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-learner --rounds 1 > $OUT/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_rollout|k_select_windows|k_make_batch|k_target_rows" -s 16 -c 4 -o $OUT/step_full python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-learner --rounds 1 > $OUT/ncu_full.log 2>&1
tail -2 $OUT/ncu_full.log
python tools/ncu_summary.py $OUT/step_full.ncu-rep > $OUT/step_full.md
# rewrites the traffic file bench.py reads; fails if a bench phase has no kernel in the capture
python tools/make_traffic.py $OUT/step_full.ncu-rep $OUT/traffic.json || exit 1
head -c 600 $OUT/bench.json
