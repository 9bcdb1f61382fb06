mkdir -p gpurun_out
for w in 192 48; do
SSF_BENCH_KL_WARM=$w timeout 600 python bench.py --workload skalow --no-e2e --no-cpu-baseline > gpurun_out/r02y_skalow_$w.json 2> gpurun_out/r02y_skalow_$w.err; tail -2 gpurun_out/r02y_skalow_$w.err
python - <<PY
import json
d=json.load(open('gpurun_out/r02y_skalow_$w.json'))
print($w, d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['alone']['frac'], d['roofline']['kernel_share_of_step'], d['fit_breakdown']['pattern_slots_in_use'])
PY
done
