#!/bin/bash
# round 2, GPU call B: full GPU suite (no -x), weight-path overlap A/B, reference arm, DRAM traffic of one internal block forward
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/b_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_tests.log
tail -15 gpurun_out/b_tests.log
python tools/ab_window.py > gpurun_out/b_ab_base.json 2> gpurun_out/b_ab_base.err
MAKANI_B200_OVERLAP_WEIGHTS=1 SFNO_WPACK_TL=32 python tools/ab_window.py > gpurun_out/b_ab_overlap32.json 2> gpurun_out/b_ab_overlap32.err
MAKANI_B200_OVERLAP_WEIGHTS=1 python tools/ab_window.py > gpurun_out/b_ab_overlap64.json 2> gpurun_out/b_ab_overlap64.err
for f in base overlap32 overlap64; do python -c "import json;d=json.load(open('gpurun_out/b_ab_$f.json'));print('$f','step',round(d['step_ms'],3),'fwd',round(d['fwd_chain_ms'],3),'blk_us',round(d['internal_block_fwd_us'],1))"; done
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/b_reference.json 2> gpurun_out/b_reference.err; echo "reference rc=$?"; cut -c1-400 gpurun_out/b_reference.json
for mb in 0 60; do
  MAKANI_B200_WINDOW_MB=$mb python tools/prof_block.py internal 3 fwd > gpurun_out/b_plain_$mb.log 2>&1 &&
  MAKANI_B200_WINDOW_MB=$mb ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
     -k regex:"kernel" --csv --log-file gpurun_out/b_traffic_internal_fwd_w$mb.csv python tools/prof_block.py internal 3 fwd > gpurun_out/b_ncu_$mb.log 2>&1
done
tail -3 gpurun_out/b_ncu_0.log; wc -l gpurun_out/b_traffic_internal_fwd_w*.csv
