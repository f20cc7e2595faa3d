# A/B of an env knob: bench without and with it (device-resident step), after the GEMM parity tests with it on
mkdir -p gpurun_out
KNOB=${1:-SFNO_GEMM_PAIR=1}
env $KNOB timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_modules.py -m gpu -q -x > gpurun_out/ab2_tests.log 2>&1; echo "tests($KNOB) rc=$?"; tail -2 gpurun_out/ab2_tests.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_ab2_off.json 2> gpurun_out/bench_ab2_off.err; echo "off rc=$?"
env $KNOB timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_ab2_on.json 2> gpurun_out/bench_ab2_on.err; echo "on rc=$?"
for f in off on; do python tools/show_bench.py gpurun_out/bench_ab2_$f.json > gpurun_out/bench_ab2_$f.txt 2>&1; echo "== $f"; cat gpurun_out/bench_ab2_$f.txt; done
