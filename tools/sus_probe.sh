# A/B of the sustained (back-to-back) figure against the per-step one under the launch / blur knobs
for cfg in "" "PL_NO_PDL=1" "PL_BLUR_FORM=reg" "PL_NO_PDL=1 PL_BLUR_FORM=reg"; do
  echo "== $cfg"
  env $cfg python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('step %.4f ms  sustained %.4f ms  e2e %.1f serial %.1f' % (d['ms_per_step'], d['sustained']['ms_per_step'], d['e2e']['value'], d['e2e'].get('serial_value',0)))
"
done
