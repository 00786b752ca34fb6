#!/bin/bash
# two GPUs: the several-GPU tests (one process, two GPUs), the checked run with --multi, bench at N=2 (torchrun)
mkdir -p gpurun_out/g2; O=gpurun_out/g2
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m pytest tests/test_delivery_and_sync.py -m gpu -q > $O/pytest_multi.txt 2>&1; echo "rc=$?" >> $O/pytest_multi.txt; tail -5 $O/pytest_multi.txt
timeout 600 python tools/sanitize_run.py --checked --multi > $O/checked_multi.txt 2>&1; echo "rc=$?" >> $O/checked_multi.txt; tail -2 $O/checked_multi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?"
tail -c 600 $O/bench_n2.err
