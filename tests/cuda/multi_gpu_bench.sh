N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --workload longform --steps 30 --warmup 5 2>/dev/null | grep "^{" | tail -1 > gpurun_out/r02_longform_$N.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --no-cpu --no-fp32 --no-sustained 2>/dev/null | grep "^{" | tail -1 > gpurun_out/r02_batched_$N.json
wc -c gpurun_out/r02_longform_$N.json gpurun_out/r02_batched_$N.json
