# scratch command file for scripts/gpu_retry.sh (what a round-end check runs): GPU tests, smoke, one bench line
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -x -m gpu -q > gpurun_out/gputest.log 2>&1; echo "pytest exit $?" >> gpurun_out/gputest.log); tail -3 gpurun_out/gputest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log); tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_lofar.json 2> gpurun_out/bench.err; echo "bench exit $?"; tail -2 gpurun_out/bench.err
