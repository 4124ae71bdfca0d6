python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/r2_probe2.py crf 2>&1 | tail -5
python bench_configs.py --workload crf_train_step --steps 5 2>&1 | tail -2
python bench_configs.py --workload crf_tsukuba --steps 5 --frames 8 2>&1 | tail -1 | cut -c1-300
