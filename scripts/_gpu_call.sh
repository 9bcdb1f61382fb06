mkdir -p gpurun_out
L=ska_sdp_screen_fitting_b200/libssf_b200.so
cp $L /tmp/lib.bak
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-extras"
$B > gpurun_out/u_lofar.json 2> gpurun_out/u_lofar.err
python - <<PY
import json
d=json.loads(open('gpurun_out/u_lofar.json').read().strip().splitlines()[-1])
print('lofar (FT=64, explicit unroll 4): step ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4), 'fit alone ms', round(d['fit']['ms_per_call_alone'],4), 'fit_pass in step ms', round(d['fit_breakdown']['kl_fit_pass']['mean_ms'],3))
PY
for u in 16 8 4; do
  if [ $u != 16 ]; then cp scripts/_libssf_unroll$u.so $L; else cp /tmp/lib.bak $L; fi
  python bench.py --workload skalow --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-extras > gpurun_out/u_skalow_$u.json 2> gpurun_out/u_skalow_$u.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/u_skalow_$u.json').read().strip().splitlines()[-1])
print('skalow unroll $u: step ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4), 'alone', round(d['roofline'].get('alone',{}).get('frac',0),4), 'fit alone ms', round(d['fit']['ms_per_call_alone'],4), d.get('fit_breakdown',{}).get('kl_fit_pass'))
PY
done
cp /tmp/lib.bak $L
