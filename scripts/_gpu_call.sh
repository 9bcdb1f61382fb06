mkdir -p gpurun_out
for a in 6 2; do
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
SSF_BENCH_FIT_AHEAD=$a timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-extras > gpurun_out/st.json 2> gpurun_out/st.err
python - <<PY
import json
d=json.load(open('gpurun_out/st.json'))
print('ahead $a', round(d['ms_per_step'],3), round(d['fit_host_ms_per_slab'],2), round(d['roofline']['frac'],3))
PY
done
done
