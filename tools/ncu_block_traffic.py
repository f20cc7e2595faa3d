"""DRAM traffic of ONE internal SpectralConv block forward from an ncu metrics capture taken with the L2 LEFT WARM
(`--cache-control none`: what one kernel leaves in L2 is there for the next, as in the real step):

    ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
        -k regex:kernel --csv --log-file X.csv python tools/prof_block.py internal 3 fwd
    python tools/ncu_block_traffic.py X.csv [label]     -> table on stdout, `block_internal_fwd` in profiles/ncu_traffic.json

The capture holds 3 forward passes; the LAST one is reported (the first ones include one-off set-up kernels).  The
weight re-pack (`pack_weight_kernel`, a layout change that training repeats every step) is listed and excluded from the
block total, as in bench.py's `roofline.block`."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "ncu_traffic.json")
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6,
         "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}


def load(path):
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.DictReader(lines)
    launches = {}
    for r in rd:
        k = launches.setdefault(r["ID"], {"name": r["Kernel Name"], "grid": r.get("Grid Size")})
        k[r["Metric Name"]] = float(r["Metric Value"].replace(",", "")) * SCALE.get(r["Metric Unit"], 1.0)
    for i in sorted(launches, key=int):
        rows.append(launches[i])
    return rows


def main():
    path = sys.argv[1]
    label = sys.argv[2] if len(sys.argv) > 2 else os.path.basename(path)
    rows = load(path)
    ours = [r for r in rows if any(s in r["name"] for s in ("dft_", "gemm_tc", "pack_weight", "xm_add_dc"))]
    # split into passes at every pack_weight launch (first kernel after the forward DFT/analysis of each pass): a pass
    # starts at the forward DFT that precedes it
    starts = [i for i, r in enumerate(ours) if "dft_fwd" in r["name"] and (i == 0 or "dft_fwd" not in ours[i - 1]["name"])
              and not any("dft_fwd" in q["name"] for q in ours[max(0, i - 1):i])]
    # robust: a pass = from a dft_fwd launch that follows a dft_inv launch (or the start) up to the next such launch
    starts = [i for i, r in enumerate(ours) if "dft_fwd" in r["name"] and (i == 0 or "dft_inv" in ours[i - 1]["name"])]
    last = ours[starts[-1]:]
    tot = 0.0
    print(f"# {label}: last of {len(starts)} forward passes, L2 left warm between kernels (ncu --cache-control none)")
    print("kernel | grid | us | dram_rd_MB | dram_wr_MB")
    for r in last:
        rd, wr = r.get("dram__bytes_read.sum", 0.0), r.get("dram__bytes_write.sum", 0.0)
        name = r["name"].split("(")[0][:48]
        print(f"{name} | {r['grid']} | {r.get('gpu__time_duration.sum', 0):.1f} | {rd / 1e6:.1f} | {wr / 1e6:.1f}")
        if "pack_weight" not in r["name"]:
            tot += rd + wr
    alg = 571115520
    print(f"block forward without the weight re-pack: {tot / 1e6:.1f} MB DRAM traffic = {tot / alg:.2f} x the algorithmic "
          f"{alg / 1e6:.1f} MB (bf16 I/O)")
    data = json.load(open(OUT)) if os.path.exists(OUT) else {}
    data["block_internal_fwd"] = {"bytes": tot, "source": label, "launches": len(last),
                                  "method": "ncu --cache-control none, dram__bytes_read.sum + dram__bytes_write.sum, summed over "
                                            "the kernels of one forward pass (weight re-pack excluded)"}
    if len(sys.argv) > 3 and sys.argv[3] == "--write":
        json.dump(data, open(OUT, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
