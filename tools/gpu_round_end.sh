#!/bin/bash
# Round-end evidence on the GPU box (run through gpurun from the repo root): GPU tests, both bench arms, the ncu launch
# list of the bench command.  Everything lands in gpurun_out/; copy what is to be judged into profiles/.
# usage: tools/gpu_round_end.sh <tag>
T=${1:-x}
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$T.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests_$T.log; tail -2 gpurun_out/gpu_tests_$T.log
python bench.py --impl reference > gpurun_out/bench_${T}_reference_arm.json 2> gpurun_out/bench_${T}_reference_arm.err; echo "reference rc=$?"
python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$T.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-parity --no-xt32 --no-other-configs > gpurun_out/ncu_launches_$T.log 2>&1; echo "launch list rc=$?"
tail -c 400 gpurun_out/bench_$T.json
