mkdir -p gpurun_out
P="timeout 120 python scripts/prof_kernel.py"
run_set() {
  echo "##### build flags: '$1'"
  SSF_NVCC_EXTRA="$1" python -m ska_sdp_screen_fitting_b200._build > /dev/null 2>&1 || echo BUILD FAILED
  export SSF_NVCC_EXTRA="$1"
  shift
  for args in "$@"; do eval "$args" 2>&1 | grep -v "^  \|^Traceback\|^torch\.\|^Search\|^CUDA\|^For debug\|^Compile\|^$" | tail -1; done
  unset SSF_NVCC_EXTRA
}
{
run_set "-DSSF_BK_STATIC" "SSF_TC_BIGK=2 $P kl --ndir 56 --n 1024 --ncol 1500 --check" "$P kl --ndir 100 --n 2048 --ncol 512"
run_set "-DSSF_BK_LATE_RAW" "SSF_TC_BIGK=2 $P kl --ndir 56 --n 1024 --ncol 1500 --check" "$P kl --ndir 100 --n 2048 --ncol 512"
run_set "-DSSF_BK_LATE_RAW -DSSF_BK_STATIC -DSSF_BK_EPI=8" "SSF_TC_BIGK=2 $P kl --ndir 56 --n 1024 --ncol 1500 --check"
} > gpurun_out/r02s_bk.txt 2>&1
cat gpurun_out/r02s_bk.txt
python -m ska_sdp_screen_fitting_b200._build > /dev/null 2>&1
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02s_gputest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02s_gputest.log); tail -4 gpurun_out/r02s_gputest.log
timeout 600 python bench.py --workload voronoi1k --no-e2e --no-cpu-baseline > gpurun_out/r02s_v1k.json 2> gpurun_out/r02s_v1k.err; tail -2 gpurun_out/r02s_v1k.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02s_v1k.json'))
print(d['value'], d['roofline']['frac']); v=d['voronoi_direct_smoothing']; print('direct', v['value'], v['ms_per_step'], v['roofline']['frac'])
PY
