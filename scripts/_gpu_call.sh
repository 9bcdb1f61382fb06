mkdir -p gpurun_out
P="python scripts/prof_kernel.py"
{
echo "# narrow-tile KL kernel, 100 directions, 2048^2: column tiles (of 32 columns) per CTA"
for per in 8 12 16; do
echo "SSF_TC_PER=$per"
SSF_TC_PER=$per $P kl --ndir 100 --n 2048 --ncol 512 --reps 20
done
for per in 8 16 32; do
echo "SSF_TC_PER=$per (1024 columns)"
SSF_TC_PER=$per $P kl --ndir 100 --n 2048 --ncol 1024 --reps 10
done
echo "# 64 directions"
for per in 6 8 16; do
SSF_TC_PER=$per $P kl --ndir 64 --n 2048 --ncol 512 --reps 20
done
} > gpurun_out/r02y_narrow_per.txt 2>&1
cat gpurun_out/r02y_narrow_per.txt
