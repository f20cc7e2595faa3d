# quick GPU check of a kernel change: kernel + module parity tests, then the device-resident bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_modules.py -m gpu -q -x > gpurun_out/ab_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/ab_tests.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_ab.json 2> gpurun_out/bench_ab.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/bench_ab.json > gpurun_out/bench_ab.txt 2>&1; cat gpurun_out/bench_ab.txt
