mkdir -p gpurun_out/f1
for c in orr pa fgm kmeans mcpa; do
timeout 400 python bench.py --config $c --steps 20 --warmup 5 > gpurun_out/f1/$c.json 2> gpurun_out/f1/$c.err; echo "$c rc $?"; python -c "
import json; j=json.loads(open('gpurun_out/f1/$c.json').read().strip().splitlines()[-1]); e=j.get('e2e') or {}; print('$c', round(j['value']/1e6,1),'M', round(j['ms_per_step']*1e3,2),'us vs_base', round(j['vs_baseline'] or 0,2), 'e2e', round((e.get('value') or 0)/1e6,1), 'frac', e.get('fraction_of_h2d_ceiling'))"
done
