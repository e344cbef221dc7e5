cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( timeout -s KILL 200 python -m pytest tests/test_gpu_round2.py tests/test_gpu_parity.py -m gpu -x -q -k "fused_tree or small_build or bvh_structure or city" ) > gpurun_out/n_tests1.log 2>&1
rc=$?; echo "tree tests rc=$rc"; tail -3 gpurun_out/n_tests1.log
for i in 1 2; do timeout -s KILL 90 python tools/build_stages.py --reps 10; done 2>&1 | tee gpurun_out/n_stages.jsonl
