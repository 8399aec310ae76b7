#!/bin/bash
# Tuning aid: build libcrt variants (extra -D defines) here, then run the short bench on each on the GPU box.
#   tools/variants.sh build  name1:DEF1,DEF2  name2:DEF3 ...      (CPU box)
#   tools/variants.sh run [bench args]                            (GPU box; prints one summary line per variant)
set -e
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
  shift; rm -rf variants; mkdir -p variants
  for spec in "$@"; do
    name=${spec%%:*}; defs=${spec#*:}; [ "$defs" = "$spec" ] && defs=""
    python - "$name" "$defs" <<'PY'
import sys
from computeraytracer_b200.build import build_variant
name, defs = sys.argv[1], [d for d in sys.argv[2].split(",") if d]
print(build_variant(f"variants/libcrt_{name}.so", defs))
PY
  done
else
  shift
  for lib in variants/libcrt_*.so; do
    CRT_LIBCRT=$PWD/$lib python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-also "$@" 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.read().strip()
try:
    d = json.loads(l); k = d['roofline']['kernels_ms_per_step']
    print('$lib', round(d['value']), 'Mrays/s', round(d['ms_per_step'], 1), 'ms', {a: round(b, 1) for a, b in k.items()})
except Exception as e:
    print('$lib', 'FAILED', l[-300:])
"
  done
fi
