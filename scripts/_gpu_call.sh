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
run_set "-DSSF_BK_SS" "SSF_TC_BIGK=2 $P kl --ndir 56 --n 1024 --ncol 1500 --check" "SSF_TC_BIGK=2 $P kl --ndir 24 --n 1024 --ncol 1500 --check" "SSF_TC_BIGK=2 $P kl --check" "SSF_TC_BIGK=2 $P gain --check"
run_set "-DSSF_BK_SS -DSSF_BK_EPI=8" "SSF_TC_BIGK=2 $P kl --ndir 56 --n 1024 --ncol 1500 --check" "SSF_TC_BIGK=2 $P kl --ndir 24 --n 1024 --ncol 1500 --check"
run_set "" "SSF_TC_BIGK=2 $P kl --ndir 56 --n 1024 --ncol 1500 --check" "SSF_TC_BIGK=2 $P kl --ndir 24 --n 1024 --ncol 1500"
} > gpurun_out/r02r_ss.txt 2>&1
cat gpurun_out/r02r_ss.txt
python -m ska_sdp_screen_fitting_b200._build > /dev/null 2>&1
