mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -x -m gpu -q > gpurun_out/r02zd_gputest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02zd_gputest.log); tail -3 gpurun_out/r02zd_gputest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02zd_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r02zd_smoke.log); tail -2 gpurun_out/r02zd_smoke.log
P="python scripts/prof_kernel.py"
{ for f in 0 1 2 3; do $P fill --ndir 45 --n 512 --ncol 5952 --flags $f; $P fill --ndir 45 --n 2048 --ncol 372 --flags $f; done; $P fillgain --ndir 45 --n 512 --ncol 5952; $P fillgain --ndir 45 --n 2048 --ncol 372; } > gpurun_out/r02zd_fill.txt 2>&1; cat gpurun_out/r02zd_fill.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02zd_bench_reference.json 2>/dev/null; echo "ref exit $?"
timeout 900 python bench.py > gpurun_out/r02zd_bench_lofar.json 2> gpurun_out/r02zd_bench.err; echo "bench exit $?"; tail -2 gpurun_out/r02zd_bench.err
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
SSF_BENCH_KL_WARM=8 $B > gpurun_out/ncu_plain_b.log 2>&1 && SSF_BENCH_KL_WARM=8 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02zd_launches.csv $B > gpurun_out/ncu_b.log 2>&1
tail -2 gpurun_out/ncu_b.log
