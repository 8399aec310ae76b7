#!/bin/bash
# The other BASELINE.json configs on one GPU (config 3 is the default bench): one JSON line each into gpurun_out/.
TAG=${1:-r01}
O=gpurun_out
python bench.py --workload cornell --steps 8 --warmup 3 --no-also > $O/${TAG}_cfg2_cornell_1080p.json 2> $O/${TAG}_cfg2.err
python bench.py --workload mesh1m --width 3840 --height 2160 --steps 3 --warmup 3 --no-also --no-cpu-baseline > $O/${TAG}_cfg4_mesh1m_4k_1gpu.json 2> $O/${TAG}_cfg4.err
python bench.py --workload soup10m --spp-per-step 16 --steps 4 --warmup 3 --no-also > $O/${TAG}_cfg5_soup10m.json 2> $O/${TAG}_cfg5.err
for f in $O/${TAG}_cfg*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1])); r = d["roofline"]
    print(sys.argv[1], round(d["value"]), "Mrays/s", round(d["total_mrays_s"]), "total", round(d["ms_per_frame"], 3), "ms/frame", r["kernel"], round(r["frac"], 3),
          {k: round(v, 1) for k, v in r["kernels_ms_per_step"].items()}, d.get("cpu_baseline", {}).get("value"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
