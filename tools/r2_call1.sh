set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c1_gpus.txt
timeout 1500 python tools/r2_probe.py > gpurun_out/c1_probe.jsonl 2> gpurun_out/c1_probe.err
echo probe rc=$?
timeout 900 compute-sanitizer --tool memcheck --log-file gpurun_out/c1_memcheck_smoke.log python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c1_memcheck_smoke.out 2>&1
echo memcheck rc=$?
tail -5 gpurun_out/c1_memcheck_smoke.log
