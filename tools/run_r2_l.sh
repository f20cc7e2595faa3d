#!/bin/bash
# round 2, GPU call L (4 GPUs): spatial blocks at h = 4 with the SM row scatter and with the copy-engine variant of the XM exchanges
mkdir -p gpurun_out
for mode in sm ce; do
  MAKANI_B200_P2P_COPY=$mode timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29577 tools/bench_spatial_blocks.py > gpurun_out/l_spatial_$mode.json 2> gpurun_out/l_spatial_$mode.err
  echo "$mode rc=$?"; grep -E "Error|error" gpurun_out/l_spatial_$mode.err | head -5 | cut -c1-300
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/l_spatial_$mode.json') if l.startswith('{')][-1])
    print('$mode', {k:(round(v['ms_h'],3),round(v['ms_1gpu'],3),round(v['speedup'],2)) for k,v in d['blocks'].items()}, 'lat_to_chan', round(d['lat_to_chan_721']['ms'],3), round(d['lat_to_chan_721']['GBs_per_gpu'],1))
except Exception as e: print('$mode failed', e)
PY
done
