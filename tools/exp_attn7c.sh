mkdir -p gpurun_out
timeout 100 python tools/att_check.py 8x485x40 1x1370x12 3x50x2 2x257x3 2x485x6 1x128x1 > gpurun_out/att_check.log 2>&1; tail -6 gpurun_out/att_check.log
timeout 300 python tools/prof_kernels.py attention 20
