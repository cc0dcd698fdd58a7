import json, sys
d = json.load(open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/bench_q.json'))
print(d['ms_per_step'], '%.3e' % d['value'], {k: round(v['ms_per_step'], 3) for k, v in d['passes'].items() if v['ms_per_step']})
s = d['config'].get('secondary')
if s:
    print(s['ms_per_step'], {k: round(v, 4) for k, v in s['pass_ms_per_step'].items() if v})
