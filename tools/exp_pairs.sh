# gpurun -- 'bash tools/exp_pairs.sh': throughput against the number of pairs per call (is a 2 x 32 split of the 64-pair step worth it?)
mkdir -p gpurun_out
for p in 64 32 16 64 32; do
  timeout 600 python bench.py --pairs $p --fast --steps 30 --warmup 5 > gpurun_out/pairs_$p.log 2>&1
  python tools/show_bench.py gpurun_out/pairs_$p.log | head -3
done
