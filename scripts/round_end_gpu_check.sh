# Full GPU check of the tree: tests, training-step timings, smoke, the bench line, ncu launch lists of one training step.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/t_full.log 2>&1; tail -3 gpurun_out/t_full.log
timeout 200 python scripts/bench_train.py --cases c2,c3 --iters 20 > gpurun_out/bench_train.jsonl 2> gpurun_out/bench_train.err; cat gpurun_out/bench_train.jsonl
timeout 100 python scripts/train_enqueue_time.py c2,c3 > gpurun_out/train_enqueue.jsonl 2>&1; tail -2 gpurun_out/train_enqueue.jsonl
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 600 gpurun_out/bench.json
python scripts/profile_train.py --case c3 > gpurun_out/pt3.log 2>&1 && timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_train_c3.csv python scripts/profile_train.py --case c3 > gpurun_out/ncu_l3.log 2>&1
python scripts/profile_train.py --case c2 > gpurun_out/pt2.log 2>&1 && timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_train_c2.csv python scripts/profile_train.py --case c2 > gpurun_out/ncu_l2.log 2>&1
ls -la gpurun_out/launches_train_c*.csv
