set -x
nvidia-smi -L
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731 tools/peer_check.py 2>&1 | tail -5
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29732 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2a_bench_2gpu.json 2> gpurun_out/r2a_bench_2gpu.err ) 2>&1 | tail -4
tail -5 gpurun_out/r2a_bench_2gpu.err
