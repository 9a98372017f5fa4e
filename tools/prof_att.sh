mkdir -p gpurun_out
timeout 100 python tools/att_check.py 8x485x40 1x1370x12 3x50x2 2x257x3 2x485x6 > gpurun_out/att_check.log 2>&1; cat gpurun_out/att_check.log | tail -5
python tools/prof_kernels.py attention 3 > gpurun_out/plain_att.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attention -s 2 -c 1 -f -o gpurun_out/prof_att python tools/prof_kernels.py attention 3 > gpurun_out/ncu_att.log 2>&1
tail -2 gpurun_out/ncu_att.log; cat gpurun_out/plain_att.log
