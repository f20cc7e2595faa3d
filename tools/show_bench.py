import json, sys
l = [x for x in open(sys.argv[1]) if x.startswith('{')]
d = json.loads(l[-1])
print("ms/step", round(d['ms_per_step'], 3), "e2e ms", round(d['e2e']['ms_per_step'], 2), "| roofline:", d['roofline']['kernel'],
      round(d['roofline']['frac'], 3), "| cpu", d.get('cpu_baseline'))
for r in d['kernels']:
    print(f"{r['kernel']:20s} K={r['shape'].get('K','-'):>4} n={r['launches_per_step']:2d} {r['avg_us']:8.1f} us share {r['share']:.3f} frac {r['frac']:.3f}")
