# gpurun -- 'bash tools/run_configs_n1.sh': BASELINE configs 3, 4, 5 on one GPU (the N=1 points of their scaling lines)
mkdir -p gpurun_out
run() { name=$1; shift; ( time timeout 900 python bench.py "$@" ) > gpurun_out/$name.log 2>&1; echo "exit $?" >> gpurun_out/$name.log; grep '^{' gpurun_out/$name.log > gpurun_out/$name.json; tail -4 gpurun_out/$name.log | cut -c1-200; }
run r2_config3_layer11_n1 --config 3 --layer 11 --fast
run r2_config3_layer23_n1 --config 3 --layer 23 --fast
run r2_config4_layer11_n1 --config 4 --layer 11 --fast
run r2_config4_layer39_n1 --config 4 --layer 39 --fast
run r2_config5_n1 --config 5 --steps 10 --warmup 3
python tools/show_bench.py gpurun_out/r2_config3_layer11_n1.json gpurun_out/r2_config3_layer23_n1.json gpurun_out/r2_config4_layer11_n1.json gpurun_out/r2_config4_layer39_n1.json
