#!/bin/bash
# like ab_libs.sh, with environment settings per run: tools/ab_env.sh "LIB [VAR=VALUE ...]" ...
for spec in "$@"; do
  set -- $spec; lib=$1; shift
  env "$@" IPKGPU_LIB=$PWD/$lib python bench.py --no-cpu-baseline --no-file --steps 3 --warmup 2 > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err || { echo "$spec FAILED"; tail -2 gpurun_out/ab_tmp.err; continue; }
  python - "$spec" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/ab_tmp.json").read().strip().splitlines()[-1])
k = {x["kernel"]: round(x["ms"] / d["steps"], 2) for x in d["kernels"]}
print(sys.argv[1], "step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["ms_per_step"], 1), {n: k.get(n) for n in ("k2_join", "k2_join_retry1", "ps_msd_pack", "ps_sweep", "mg_merge_packed", "k4_mif0")})
PY
done
