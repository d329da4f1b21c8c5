#!/bin/bash
# A/B of library variants built by csrc/build.py build_variant: per-stage times of the cfg3 step for each
for v in "$@"; do
  ASB_LIB_PATH=$PWD/activesplat_b200/variants/lib_$v.so python bench.py --extras none --no-ref-cuda --no-cpu-baseline --steps 30 > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    r = json.load(open(f"gpurun_out/ab_{v}.json"))
    print(v, "graph %.4f eager %.4f" % (r["ms_per_step"], r["eager"]["ms_per_step"]), {k: s["ms"] for k, s in r["stages"].items()}, "loss", r["final_loss"])
except Exception as ex:
    print(v, "FAILED", ex, open(f"gpurun_out/ab_{v}.err").read()[-600:])
PY
done
