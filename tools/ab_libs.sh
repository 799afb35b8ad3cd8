#!/bin/bash
# A/B of library builds on one box: tools/ab_libs.sh libA.so libB.so ...   (IPKGPU_LIB picks the build, ipk_b200/capi.py)
for lib in "$@"; do
  IPKGPU_LIB=$PWD/$lib python bench.py --no-cpu-baseline --no-file --steps 3 --warmup 2 > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err || { echo "$lib FAILED"; tail -2 gpurun_out/ab_tmp.err; continue; }
  python - "$lib" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/ab_tmp.json").read().strip().splitlines()[-1])
k = {x["kernel"]: round(x["ms"] / d["steps"], 2) for x in d["kernels"]}
print(sys.argv[1], "step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["ms_per_step"], 1), {n: k.get(n) for n in ("k2_join", "ps_msd_pack", "ps_sweep", "mg_merge_packed", "k4_mif0")})
PY
done
