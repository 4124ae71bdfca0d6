G=$1; TAG=$2
python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29752 bench.py --gpus $G --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_${G}gpu.json 2> gpurun_out/${TAG}_bench_${G}gpu.err
tail -2 gpurun_out/${TAG}_bench_${G}gpu.err
