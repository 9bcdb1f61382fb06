mkdir -p gpurun_out
P="timeout 120 python scripts/prof_kernel.py"
{
$P kl --check
$P gain --check
$P kl --flags 3 
$P voronoi --n 1024 --ncol 1488
$P voronoi --n 1024 --ncol 1488 --flags 3
$P kl --ndir 100 --n 2048 --ncol 512 --check
$P gain --ndir 100 --n 2048 --ncol 512 --check
$P voronoi --ndir 100 --n 2048 --ncol 512
SSF_TC_BIGK=2 $P gain --check
SSF_TC_BIGK=2 $P kl --check
$P fill --ndir 100 --n 2048 --ncol 512
$P fill --ndir 45 --n 512 --ncol 5952
$P kl --ndir 60 --n 1024 --ncol 1500 --check
$P gain --ndir 60 --n 1024 --ncol 1500 --check
$P gain --ndir 112 --n 1024 --ncol 1500 --check
$P gain --ndir 9 --n 1024 --ncol 1500 --check
$P kl --ndir 9 --n 1024 --ncol 1500 --check
} > gpurun_out/r02d_prof.txt 2>&1
grep -v "^  \|^Traceback\|^torch\.\|^Search\|^CUDA\|^For debug\|^Compile\|^$" gpurun_out/r02d_prof.txt
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02d_gputest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02d_gputest.log); tail -8 gpurun_out/r02d_gputest.log
