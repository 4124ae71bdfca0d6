TAG=$1
python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_ref.err
python bench_configs.py --workload crf_train_step --steps 5 > gpurun_out/${TAG}_train.jsonl 2>/dev/null
tail -2 gpurun_out/${TAG}_train.jsonl | cut -c1-160
