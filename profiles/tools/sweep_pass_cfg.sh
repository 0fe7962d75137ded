#!/bin/bash
# bench.py once per radix pass configuration: "<cfg> <ms/step> <avg pass ms>"
for c in "$@"; do
  python bench.py --steps 8 --warmup 3 --no-e2e --no-cpu-baseline --radix-config "$c" 2>/dev/null | tail -1 > /tmp/sweep_line.json
  python - "$c" <<'PY'
import json, sys
j = json.load(open("/tmp/sweep_line.json"))
r = j.get("roofline_second_kernel") or j.get("roofline")
print(sys.argv[1], round(j["ms_per_step"], 3), round(r["avg_launch_ms"], 4), r["kernel"][:30])
PY
done
