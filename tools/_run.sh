cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( timeout -s KILL 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "small_build or golden or transform_bit_exact or strided" ) > gpurun_out/m_tests1.log 2>&1
rc=$?; echo "small tests rc=$rc"; tail -3 gpurun_out/m_tests1.log
if [ $rc -ne 0 ]; then exit 1; fi
( timeout -s KILL 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/m_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/m_tests.log
timeout -s KILL 300 python bench.py --no-cpu-baseline --no-parity > gpurun_out/m_bench.json 2>/dev/null
python -c "
import json; b=json.load(open('gpurun_out/m_bench.json')); print(b['value']); print(b['xt32_scan']); print(b['other_configs']['c2'])"
timeout -s KILL 100 python tools/small_scan_probe.py 2>&1 | tail -3
