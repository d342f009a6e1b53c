run() { TAG="$1" env $2 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $3 profiles/nccl_probe.py 2>&1 | grep -E "all_reduce|rror" ; }
run default "X=1" 29531
run nvls "NCCL_ALGO=NVLS" 29532
run tree "NCCL_ALGO=Tree" 29533
NCCL_DEBUG=INFO python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 profiles/nccl_probe.py 2>&1 | grep -iE "nvls|algo|channels" | head -8
