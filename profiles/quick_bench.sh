python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/q.json 2> gpurun_out/q.err || tail -5 gpurun_out/q.err
python -c "
import json; d=json.load(open('gpurun_out/q.json')); print('value %.0f step %.2f bwd %.2f ms (%.3f) fwd %.2f ms (%.3f)' % (d['value'], d['ms_per_step'], d['roofline']['ms_per_launch'], d['roofline']['frac'], d['roofline_fwd']['ms_per_launch'], d['roofline_fwd']['frac']), d['clocks'])"
