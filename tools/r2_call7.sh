python tools/r2_one.py 500 740 64 sharp 3 || exit 1
ncu --set full --clock-control none --import-source on -k regex:blur_all_kernel -s 6 -c 1 -o gpurun_out/r2b_blur_sharp -f python tools/r2_one.py 500 740 64 sharp 3 > gpurun_out/r2b_ncu_blur.log 2>&1
tail -3 gpurun_out/r2b_ncu_blur.log
ncu --set full --clock-control none --import-source on -k regex:blur_all_kernel -s 6 -c 1 -o gpurun_out/r2b_blur_textured -f python tools/r2_one.py 500 740 64 sharp_textured 3 > gpurun_out/r2b_ncu_blur2.log 2>&1
tail -3 gpurun_out/r2b_ncu_blur2.log
python tools/r2_probe2.py crf 2>&1 | tail -6
