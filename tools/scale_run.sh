#!/bin/bash
# bench.py at N GPUs (torchrun, one rank per GPU): C4 headline + C5.   tools/scale_run.sh N
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 --no-c2 > gpurun_out/r02_final_n${N}.json 2> gpurun_out/r02_final_n${N}.err; echo "c4 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload c5 --steps 1 --warmup 1 --e2e-steps 1 --no-c2 > gpurun_out/r02_final_n${N}_c5.json 2> gpurun_out/r02_final_n${N}_c5.err; echo "c5 rc=$?"
