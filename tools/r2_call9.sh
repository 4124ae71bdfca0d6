G=$1
nvidia-smi -L | wc -l
python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29741 tools/peer_check.py 2>&1 | tail -2
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29742 bench.py --gpus $G --steps 20 --warmup 5 > gpurun_out/r2c_bench_${G}gpu.json 2> gpurun_out/r2c_bench_${G}gpu.err ) 2>&1 | tail -4
tail -3 gpurun_out/r2c_bench_${G}gpu.err
