# tile shapes of the one-sweep radix pass on the default workload (run under gpurun)
for sh in ${SWEEP_SHAPES:-2 6 7 8 9 10 11}; do
  IPKGPU_RS_SHAPE=$sh python bench.py --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('rs_shape $sh', round(d['ms_per_step'],2), d['stages_ms_rank0']['ms_sort'])"
done
