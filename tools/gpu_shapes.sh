#!/bin/bash
# The other BASELINE.json data shapes (zeros / random / text-like), device-resident + e2e; one summary line each.
mkdir -p gpurun_out
for w in ${SHAPES:-zeros random textlike loglike}; do
  python bench.py --steps 5 --warmup 3 --skip-cpu --workload $w ${BENCH_ARGS} > gpurun_out/shape_$w.json 2> gpurun_out/shape_$w.err || tail -3 gpurun_out/shape_$w.err
  python - "$w" <<'PY'
import json, sys
w = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/shape_{w}.json").read().strip().splitlines()[-1])
    r, o = d["roofline"], d["roofline_other"]
    print(f"{w:9s} compress {d['compress_GBps']:8.1f} GB/s  decompress {d['decompress_GBps']:8.1f} GB/s  ratio {d['ratio']:8.3f}  "
          f"{r['kernel']} {r['frac']:.3f}  {o['kernel']} {o['frac']:.3f}  e2e {d['e2e']['value'] if d['e2e'] else 0:.1f}")
except Exception as e:
    print(w, "failed", e)
PY
done
