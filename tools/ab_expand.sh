run() {  # $1 label
  timeout 200 python bench.py --steps 50 --warmup 5 --repeats 7 --no-cpu-baseline --no-churn 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'step %.2f [%.2f %.2f]' % (1e3*d['ms_per_step_spread']['median'], 1e3*d['ms_per_step_spread']['min'], 1e3*d['ms_per_step_spread']['max']), 'expand %.2f prep %.2f' % (1e3*d['kernel_ms']['expand'], 1e3*d['kernel_ms']['prep']), 'frac %.3f alone %.3f step_frac %.3f' % (d['roofline']['frac'], d['roofline']['frac_alone'], d['roofline']['step_frac']))"
}
for v in $VARIANTS; do
  export GS_VARIANT=$v
  [ "$v" = base ] && export GS_VARIANT=""
  timeout 300 python -m pytest tests/test_gpu_filter.py -m gpu -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -1
  run "variant=[$v]"
done
