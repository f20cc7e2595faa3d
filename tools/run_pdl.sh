mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_modules.py -m gpu -q -x > gpurun_out/pdl_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/pdl_tests.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_pdl_on.json 2> gpurun_out/bench_pdl_on.err; echo "on rc=$?"
SFNO_PDL=0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_pdl_off.json 2> gpurun_out/bench_pdl_off.err; echo "off rc=$?"
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_pdl_on2.json 2> gpurun_out/bench_pdl_on2.err; echo "on2 rc=$?"
for f in on off on2; do python tools/show_bench.py gpurun_out/bench_pdl_$f.json | head -1; done
timeout 300 python tools/prof_block.py block7 1 > gpurun_out/plain_b7.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"dft_inv_p2_kernel|gemm_tc_kernel" -c 6 -o gpurun_out/r1f_full_block7 -f \
    python tools/prof_block.py block7 1 > gpurun_out/ncu_f_block7.log 2>&1; echo "ncu rc=$?"
