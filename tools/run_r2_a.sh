#!/bin/bash
# round 2, GPU call A: tests + smoke + window A/B + bench + DRAM traffic of one internal block forward (ncu, L2 left warm)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/a_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_tests.log
tail -5 gpurun_out/a_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/a_smoke.log 2>&1; tail -2 gpurun_out/a_smoke.log
for mb in 0 30 60 90 120; do
  MAKANI_B200_WINDOW_MB=$mb python tools/ab_window.py > gpurun_out/a_ab_w${mb}.json 2> gpurun_out/a_ab_w${mb}.err
  python -c "import json;d=json.load(open('gpurun_out/a_ab_w${mb}.json'));print('window',d['window_mb'],'step',round(d['step_ms'],3),'fwd',round(d['fwd_chain_ms'],3),'blk_us',round(d['internal_block_fwd_us'],1))"
done
MAKANI_B200_WINDOW_MB=60 SFNO_WINDOW_HINTS=0 python tools/ab_window.py > gpurun_out/a_ab_w60_nohints.json 2> gpurun_out/a_ab_w60_nohints.err
python -c "import json;d=json.load(open('gpurun_out/a_ab_w60_nohints.json'));print('window 60 nohints step',round(d['step_ms'],3),'fwd',round(d['fwd_chain_ms'],3),'blk_us',round(d['internal_block_fwd_us'],1))"
python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/a_bench.err
python -c "
import json;d=json.load(open('gpurun_out/a_bench.json'));print('ms/step',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'block',d['roofline'].get('block',{}).get('us'),'gpu_base',(d.get('gpu_baseline') or {}).get('ms_per_step'),'rollout',(d.get('rollout') or {}).get('ms_per_step_eager'),(d.get('rollout') or {}).get('ms_per_step_graph'))"
for mb in 0 60; do
  MAKANI_B200_WINDOW_MB=$mb python tools/prof_block.py internal 3 fwd > gpurun_out/a_plain_$mb.log 2>&1 &&
  MAKANI_B200_WINDOW_MB=$mb ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
     -k regex:"sfno" --csv --log-file gpurun_out/a_traffic_internal_fwd_w$mb.csv python tools/prof_block.py internal 3 fwd > gpurun_out/a_ncu_$mb.log 2>&1
done
ls -la gpurun_out | head -40
