# gpurun -- 'bash tools/profile_round.sh': ncu launch list of the bench command (set-up, warm-up and the timed device-resident
# steps) and one --set full capture of the hot kernels, each only after the same command has exited 0 without ncu.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --fast"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 600 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|attention_ts|descriptors_unit|im2col_kernel" -s 119 -c 14 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_launches.log; tail -2 gpurun_out/ncu_full.log; ls -la gpurun_out/prof.ncu-rep gpurun_out/launches.csv
