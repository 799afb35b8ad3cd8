# tuning sweep of the join kernel's knobs on the default workload (run under gpurun)
run() { python bench.py --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],2), d['stages_ms_rank0']['ms_enumerate'])"; }
for gm in 5 8 10 14; do for sl in 32; do IPKGPU_GROUP_MIN=$gm IPKGPU_SMALL_LIST=$sl run "G4 group_min $gm small_list $sl"; done; done
export IPKGPU_LIB=$PWD/ipk_b200/libipkgpu_g8.so
for gm in 8 12 20; do for sl in 32 64; do IPKGPU_GROUP_MIN=$gm IPKGPU_SMALL_LIST=$sl run "G8 group_min $gm small_list $sl"; done; done
