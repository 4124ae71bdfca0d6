M="lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,sm__inst_executed_pipe_tensor.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active,l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum,l1tex__t_set_accesses_pipe_lsu_mem_global_op_atom.sum"
python bench.py --only-step --steps 3 --warmup 3 > gpurun_out/r2d_only.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2d_launches.csv python bench.py --only-step --steps 3 --warmup 3 > gpurun_out/r2d_ncu1.log 2>&1
ncu --set full --metrics $M --clock-control none --import-source on -k regex:mf_ws_kernel -s 21 -c 1 -o gpurun_out/r2d_ws -f python bench.py --only-step --steps 3 --warmup 3 > gpurun_out/r2d_ncu2.log 2>&1
tail -2 gpurun_out/r2d_ncu2.log
ncu --set full --metrics $M --clock-control none --import-source on -k "regex:splat_pipe|compat_tc_kernel|slice_tiled" -s 15 -c 3 -o gpurun_out/r2d_L256 -f python tools/r2_one.py 1450 2000 256 default 2 > gpurun_out/r2d_ncu3.log 2>&1
tail -2 gpurun_out/r2d_ncu3.log
ncu --set full --metrics $M --clock-control none -k "regex:splat_sorted|splat_agg" -c 2 -o gpurun_out/r2d_smallc -f python -c "
import sys; sys.path.insert(0,'.')
import torch, depth_estimation_b200 as pkg, bench_util as U
f=torch.from_numpy(U.features(1024,1024,'sharp')).cuda(); lat=pkg.PermutohedralLattice(f)
for C in (4,16):
    s=torch.rand((1024*1024,C),device='cuda'); lat.filter(s)
torch.cuda.synchronize()
" > gpurun_out/r2d_ncu4.log 2>&1
tail -2 gpurun_out/r2d_ncu4.log
