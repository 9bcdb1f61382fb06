mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02ab_gputest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02ab_gputest.log); tail -3 gpurun_out/r02ab_gputest.log
for i in 1 2 3; do
timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-extras > gpurun_out/st.json 2> gpurun_out/st.err; tail -1 gpurun_out/st.err
python - <<PY
import json
d=json.load(open('gpurun_out/st.json'))
print('lofar', d['value'], round(d['ms_per_step'],3), round(d['roofline']['frac'],3), round(d['roofline']['alone']['frac'],3), round(d['fit_host_ms_per_slab'],2), d['fit']['ms_per_call_alone'], {k:round(v['mean_ms'],2) for k,v in d['fit_breakdown'].items() if isinstance(v,dict) and 'mean_ms' in v})
PY
done
for i in 1 2; do
timeout 600 python bench.py --workload skalow --no-e2e --no-cpu-baseline > gpurun_out/st.json 2> gpurun_out/st.err; tail -1 gpurun_out/st.err
python - <<PY
import json
d=json.load(open('gpurun_out/st.json'))
print('skalow', d['value'], round(d['ms_per_step'],3), round(d['roofline']['frac'],3), round(d['roofline']['alone']['frac'],3), round(d['fit_host_ms_per_slab'],2), d['fit']['ms_per_call_alone'])
PY
done
