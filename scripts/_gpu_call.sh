mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02z_gputest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02z_gputest.log); tail -3 gpurun_out/r02z_gputest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02z_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r02z_smoke.log); tail -2 gpurun_out/r02z_smoke.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02z_bench_reference.json 2>/dev/null; echo "ref exit $?"
timeout 900 python bench.py > gpurun_out/r02z_bench_lofar.json 2> gpurun_out/r02z_bench.err; echo "bench exit $?"; tail -2 gpurun_out/r02z_bench.err
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
SSF_BENCH_KL_WARM=8 $B > gpurun_out/ncu_plain_b.log 2>&1 && SSF_BENCH_KL_WARM=8 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02z_launches.csv $B > gpurun_out/ncu_b.log 2>&1
tail -2 gpurun_out/ncu_b.log
P="python scripts/prof_kernel.py"
$P kl --ndir 100 --n 2048 --ncol 512 --reps 1 --warmup 1 > gpurun_out/p3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:contract_tc -c 1 -o gpurun_out/r02z_kl100 -f $P kl --ndir 100 --n 2048 --ncol 512 --reps 1 --warmup 1 > gpurun_out/n3.log 2>&1
tail -1 gpurun_out/n3.log
