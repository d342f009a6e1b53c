# Config B (BASELINE.json configs[2]): 3 M Gaussians, 64 cameras at 1080p sharded over N GPUs, one gradient all-reduce
# per step.  usage: gpurun --gpus N -- 'bash profiles/gpu_b.sh N'
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --config B --total-views 64 --steps 5 --warmup 3 > gpurun_out/bench_B_n$N.json 2> gpurun_out/bench_B_n$N.err; echo rc=$?
tail -3 gpurun_out/bench_B_n$N.err | cut -c1-300; python profiles/show_bench.py gpurun_out/bench_B_n$N.json
