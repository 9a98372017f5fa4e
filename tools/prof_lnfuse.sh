# gpurun -- 'bash tools/prof_lnfuse.sh': ncu --set full of the folded-LayerNorm GEMMs in the steady part of one step
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --fast"
timeout 600 $CMD > gpurun_out/plain_lnf.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc" -s 60 -c 4 -f -o gpurun_out/prof_lnf $CMD > gpurun_out/ncu_lnf.log 2>&1
tail -2 gpurun_out/ncu_lnf.log
ncu -i gpurun_out/prof_lnf.ncu-rep --page raw --csv > gpurun_out/prof_lnf_raw.csv 2>/dev/null
for k in 0 1 2 3; do
ncu -i gpurun_out/prof_lnf.ncu-rep --page source --print-source cuda --csv --launch-skip $k --launch-count 1 > gpurun_out/prof_lnf_cuda$k.csv 2>/dev/null
done
ls -la gpurun_out/prof_lnf*
