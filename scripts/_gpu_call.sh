mkdir -p gpurun_out
P="timeout 120 python scripts/prof_kernel.py"
{
SSF_TC_BIGK=0 SSF_TC_PER=24 $P kl --ndir 57 --n 1024 --ncol 1500
SSF_TC_BIGK=0 $P kl --ndir 100 --n 2048 --ncol 512 --check
SSF_TC_BIGK=0 SSF_TC_PER=4 $P kl --ndir 100 --n 2048 --ncol 512
SSF_TC_BIGK=0 $P gain --ndir 100 --n 2048 --ncol 512 --check
SSF_TC_BIGK=0 $P voronoi --ndir 100 --n 2048 --ncol 512
SSF_TC_BIGK=0 $P kl --ndir 100 --n 1024 --ncol 2048
SSF_TC_BIGK=0 SSF_TC_PER=32 $P kl --ndir 100 --n 1024 --ncol 2048
$P kl --ndir 100 --n 1024 --ncol 2048
} > gpurun_out/r02o_cmp.txt 2>&1
grep -v "^  \|^Traceback\|^torch\.\|^Search\|^CUDA\|^For debug\|^Compile\|^$" gpurun_out/r02o_cmp.txt
