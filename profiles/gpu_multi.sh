# One 8-GPU box: config B (3 M Gaussians, 64 cameras, strong scaling) at 2 / 4 / 8 GPUs and config A at 4 GPUs.
for n in 2 4 8; do bash profiles/gpu_b.sh $n; done
bash profiles/gpu_n8.sh 4
