cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( timeout -s KILL 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sort or bvh_structure or small_build" ) > gpurun_out/r11_tests.log 2>&1
echo "sort tests rc=$?"; tail -2 gpurun_out/r11_tests.log
timeout -s KILL 90 python tools/build_stages.py --reps 10 2>&1 | tee gpurun_out/r11_stages.jsonl
for B in 1 8; do LSB200_LIB=$PWD/build_ab/liblsb200_look$B.so timeout -s KILL 90 python tools/build_stages.py --reps 10; done 2>&1 | tee -a gpurun_out/r11_stages.jsonl
