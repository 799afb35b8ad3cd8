for lv in "4000:128,20000:128,0:128,0:16,0:1" "6000:128,20000:128,0:128,0:16,0:1" "8000:128,20000:128,0:128,0:16,0:1" "12000:128,0:128,0:16,0:1" "6000:128,12000:128,0:128,0:16,0:1"; do
  IPK_BENCH_WORKLOAD=cfg4_aa_2k IPKGPU_LEVELS=$lv python bench.py --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$lv', round(d['ms_per_step'],2), d['stages_ms_rank0']['ms_enumerate'])"
done
