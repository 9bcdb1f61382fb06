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
run_set "-DSSF_NW_WRAP=65536" "$P kl --ndir 100 --n 2048 --ncol 512" "$P kl --ndir 100 --n 1024 --ncol 2048"
run_set "" "$P kl --ndir 100 --n 2048 --ncol 512" "$P kl --ndir 100 --n 1024 --ncol 2048" "$P kl --ndir 100 --n 2048 --ncol 256" "$P kl --ndir 100 --n 2048 --ncol 1024" "SSF_TC_NARROW=2 $P kl --ndir 45 --n 2048 --ncol 372" "$P kl --ndir 45 --n 2048 --ncol 372" "$P kl --ndir 64 --n 2048 --ncol 512" "$P kl --ndir 80 --n 2048 --ncol 512"
} > gpurun_out/r02v_wrap.txt 2>&1
cat gpurun_out/r02v_wrap.txt
python -m ska_sdp_screen_fitting_b200._build > /dev/null 2>&1
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
