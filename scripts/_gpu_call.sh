mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02y_noalloc_alone.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02y_noalloc_alone.log); tail -4 gpurun_out/r02y_noalloc_alone.log
(timeout 1200 python -m pytest tests -x -m gpu -q > gpurun_out/r02y_gputest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02y_gputest.log); tail -4 gpurun_out/r02y_gputest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02y_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r02y_smoke.log); tail -2 gpurun_out/r02y_smoke.log
