mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -x -m gpu -q > gpurun_out/r02zc_gputest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02zc_gputest.log); tail -3 gpurun_out/r02zc_gputest.log
timeout 900 python bench.py > gpurun_out/r02zc_bench_lofar.json 2> gpurun_out/r02zc_bench.err; echo "bench exit $?"; tail -2 gpurun_out/r02zc_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02zc_bench_lofar.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'frac',d['roofline']['frac'])
print({k:v for k,v in d.items() if 'fit' in k})
for k,v in d.items():
    if isinstance(v,dict) and 'roofline' in v: print(k, v.get('value'), v.get('ms_per_step'), v['roofline'].get('frac'))
for k,v in d.get('extra',{}).items(): print(k, v['value'], v['ms_per_step'], v['roofline']['frac'], {a:b for a,b in v.items() if 'fit' in a})
e=d['e2e']; print('e2e',e['value'],'without_file',e['without_file']['value'])
PY
