"""Collect per-kernel DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum, per launch) from ncu --set full
reports into profiles/ncu_traffic.json, keyed the way bench.py names kernels.
   python tools/ncu_traffic.py gpurun_out/a.ncu-rep gpurun_out/b.ncu-rep ..."""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "ncu_traffic.json")
NAMES = {"gemm_tc_kernel<0>": "legendre_analysis", "gemm_tc_kernel<1>": "legendre_synthesis",
         "gemm_tc_kernel<2>": "dhconv_fwd", "gemm_tc_kernel<3>": "dhconv_bwd_x",
         "gemm_tc_kernel<4>": "dhconv_bwd_w", "pack_weight_kernel": "pack_weight",
         "unpack_wgrad_kernel": "unpack_wgrad", "dft_fwd_p2_kernel<1,": "dft_fwd_480",
         "dft_fwd_p2_kernel<3,": "dft_fwd_1440", "dft_inv_p2_kernel<1,": "dft_inv_480",
         "dft_inv_p2_kernel<3,": "dft_inv_1440"}
data = json.load(open(OUT)) if os.path.exists(OUT) else {}
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, u = rows[0], rows[1]
    ik, ir, iw, it = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("gpu__time_duration.sum")
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
    for r in rows[2:]:
        for pat, name in NAMES.items():
            if pat in r[ik]:
                b = float(r[ir]) * scale[u[ir]] + float(r[iw]) * scale[u[iw]]
                # block0 / block7 captures also contain 240-latitude launches: key Legendre kernels by traffic size
                key = name
                if name.startswith("legendre"):
                    key += "_721" if b > 5e8 else "_240"
                e = data.setdefault(key, {"bytes": 0.0, "n": 0, "us": 0.0, "source": []})
                e["bytes"] = (e["bytes"] * e["n"] + b) / (e["n"] + 1)
                e["us"] = (e["us"] * e["n"] + float(r[it])) / (e["n"] + 1)
                e["n"] += 1
                if os.path.basename(rep) not in e["source"]:
                    e["source"].append(os.path.basename(rep))
json.dump(data, open(OUT, "w"), indent=1, sort_keys=True)
for k, v in sorted(data.items()):
    print(f"{k:24s} {v['bytes'] / 1e6:9.1f} MB  {v['us']:8.1f} us (ncu, cold)  n={v['n']}")
