# group width / thresholds of the join kernel's phase A2 (run under gpurun; libipkgpu_g8.so = build with -DIPK_JOIN_GROUP=8)
run() { python bench.py --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],2), d['stages_ms_rank0']['ms_enumerate'])"; }
IPKGPU_GROUP_MIN=26 run "G4 gm26"
IPKGPU_GROUP_MIN=31 run "G4 gm31"
export IPKGPU_LIB=$PWD/ipk_b200/libipkgpu_g8.so
IPKGPU_GROUP_MIN=24 IPKGPU_SMALL_LIST=64 run "G8 gm24 sl64"
IPKGPU_GROUP_MIN=28 IPKGPU_SMALL_LIST=64 run "G8 gm28 sl64"
IPKGPU_GROUP_MIN=31 IPKGPU_SMALL_LIST=64 run "G8 gm31 sl64"
IPK_BENCH_WORKLOAD=cfg4_aa_2k IPKGPU_GROUP_MIN=24 IPKGPU_SMALL_LIST=64 run "cfg4 G8 gm24 sl64"
IPK_BENCH_WORKLOAD=cfg2_dna_1k IPKGPU_GROUP_MIN=24 IPKGPU_SMALL_LIST=64 run "cfg2 G8 gm24 sl64"
