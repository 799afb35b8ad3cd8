#!/bin/bash
# join-kernel levels (pool:width,...) on cfg3: tools/sweep_levels.sh  -> ms per step and per level
for lv in "" "3072:128,6144:128,12288:128,0:128,0:16,0:1" "3072:128,8192:128,0:128,0:16,0:1" "3584:128,12288:128,0:128,0:16,0:1" "3072:128,5120:128,12288:128,0:128,0:16,0:1" "3072:128,4608:128,8192:128,0:128,0:16,0:1" "3072:128,3072:64,12288:64,0:64,0:16,0:1" "3072:128,4096:64,12288:64,0:64,0:16,0:1" "3072:128,3072:64,3072:32,0:32,0:16,0:1"; do
  IPKGPU_LEVELS=$lv python bench.py --no-cpu-baseline --no-file --steps 3 --warmup 2 > gpurun_out/lv_tmp.json 2>/dev/null
  python - "$lv" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/lv_tmp.json").read().strip().splitlines()[-1])
print(repr(sys.argv[1]), round(d["ms_per_step"], 2), "enum", d["stages_ms_rank0"]["ms_enumerate"], [(k["kernel"], k["launches"] // 3, round(k["ms"] / 3, 3)) for k in d["kernels"] if k["kernel"].startswith("k2")])
PY
done
