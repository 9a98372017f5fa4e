mkdir -p gpurun_out
python tools/prof_kernels.py attention 3 > gpurun_out/plain_att.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attention -s 2 -c 1 -f -o gpurun_out/prof_att python tools/prof_kernels.py attention 3 > gpurun_out/ncu_att.log 2>&1
tail -2 gpurun_out/ncu_att.log
