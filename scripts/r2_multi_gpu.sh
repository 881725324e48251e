#!/bin/bash
# cfg5 sweep (2048 utterances, text -> wave) on N GPUs of one box, as the driver launches it.  usage: r2_multi_gpu.sh N
N=${1:-8}
O=gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $O/r2z_bench_cfg5_${N}gpu.json 2> $O/r2z_bench_cfg5_${N}gpu.err
tail -3 $O/r2z_bench_cfg5_${N}gpu.err; cut -c1-400 $O/r2z_bench_cfg5_${N}gpu.json
