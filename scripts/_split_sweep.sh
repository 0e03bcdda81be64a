timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('step ms', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'bwd ms', round(d['roofline']['kernel_ms'],3), 'frac', round(d['roofline']['frac'],4), 'fwd ms', round(d['roofline']['other_kernels']['hashgrid_axis_fwd_kernel']['ms'],3))"
