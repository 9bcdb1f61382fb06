mkdir -p gpurun_out
for v in 1 0; do
SSF_TC_NARROW=$([ $v = 1 ] && echo 1 || echo 0) timeout 600 python bench.py --workload skalow --no-e2e --no-cpu-baseline > gpurun_out/r02u_skalow_$v.json 2> gpurun_out/r02u_skalow_$v.err; echo "narrow=$v exit $?"; tail -2 gpurun_out/r02u_skalow_$v.err
python - <<PY
import json
d=json.load(open('gpurun_out/r02u_skalow_$v.json'))
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['alone']['frac'], d['roofline']['kernel_share_of_step'], d['fit'])
PY
done
timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-extras > gpurun_out/r02u_lofar.json 2> gpurun_out/r02u_lofar.err; tail -2 gpurun_out/r02u_lofar.err
python - <<PY
import json
d=json.load(open('gpurun_out/r02u_lofar.json'))
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['alone']['frac'], d['fit'])
PY
