set -x
python -m pytest tests/test_large_labels_gpu.py -m gpu -x -q 2>&1 | tail -6
( time python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err ) 2>&1 | tail -4
tail -3 gpurun_out/r2a_bench.err
( time python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2a_bench_reference.json 2> gpurun_out/r2a_ref.err ) 2>&1 | tail -4
cat gpurun_out/r2a_bench_reference.json | cut -c1-600
nproc
