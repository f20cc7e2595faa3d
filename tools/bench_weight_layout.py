"""pack_weight / unpack_wgrad alone at the named size (Cin = Cout = 384, L = 240: 283 MB each way), CUDA events, inputs
rotated over three buffers (each 283 MB > 126 MB L2).  SFNO_WLAYOUT_CTAS = CTAs per SM of the persistent unpack kernel."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from makani_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
Cin = Cout = 384
L = 240
ws = [torch.randn(Cin, Cout, L, dtype=torch.complex64, device=dev) for _ in range(3)]
wps = [ops.pack_weight(w) for w in ws]
peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]


def timed(fn, n=30):
    for i in range(6):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


nbytes = 2 * Cin * Cout * L * 8
t_pack = timed(lambda i: ops.pack_weight(ws[i % 3]))
t_unpack = timed(lambda i: ops.unpack_wgrad(wps[i % 3], Cin, Cout, L))
ok = bool(torch.equal(ops.unpack_wgrad(wps[0], Cin, Cout, L), torch.view_as_complex(
    (torch.view_as_real(ws[0]).view(torch.int32) + 0x1000 & ~0x1FFF).view(torch.float32))))
print(json.dumps({"unpack_ctas_per_sm": os.environ.get("SFNO_WLAYOUT_CTAS", "4"), "pack_us": round(t_pack, 1), "unpack_us": round(t_unpack, 1),
                  "pack_frac": round(nbytes / t_pack / 1e3 / peak, 3), "unpack_frac": round(nbytes / t_unpack / 1e3 / peak, 3),
                  "roundtrip_is_tf32_rounding": ok}))
