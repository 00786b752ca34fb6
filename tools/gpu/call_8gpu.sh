#!/bin/bash
# eight GPUs: bench at N=8 and N=4 (C3), N=8 for C4
mkdir -p gpurun_out/g8; O=gpurun_out/g8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 3 > $O/bench_n8.json 2> $O/bench_n8.err; echo "bench n8 rc=$?"
tail -c 400 $O/bench_n8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --steps 20 --warmup 3 > $O/bench_n4.json 2> $O/bench_n4.err; echo "bench n4 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3 --workload c4 > $O/bench_c4_n8.json 2> $O/bench_c4_n8.err; echo "bench c4 n8 rc=$?"
