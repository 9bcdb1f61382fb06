mkdir -p gpurun_out
P="python scripts/prof_kernel.py"
L=ska_sdp_screen_fitting_b200/libssf_b200.so
{
echo "# two-CTA KL kernel, 45 directions: product library, then the SSF_TC_WRAP=64 build (every CTA reads its basis tile from the first 64 pixel tiles: an L2-resident window, wrong results)"
$P kl --ndir 45 --n 2048 --ncol 372
$P kl --ndir 45 --n 1024 --ncol 1488
$P kl --ndir 45 --n 512 --ncol 5952
$P kl --ndir 56 --n 2048 --ncol 372
cp $L /tmp/lib.bak; cp scripts/_libssf_wrap64.so $L
echo "# SSF_TC_WRAP=64"
$P kl --ndir 45 --n 2048 --ncol 372
$P kl --ndir 45 --n 1024 --ncol 1488
$P kl --ndir 45 --n 512 --ncol 5952
$P kl --ndir 56 --n 2048 --ncol 372
cp /tmp/lib.bak $L
} > gpurun_out/r02y_tc_wrap.txt 2>&1
cat gpurun_out/r02y_tc_wrap.txt
