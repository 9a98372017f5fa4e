# gpurun --gpus N -- 'bash tools/run_configs_multi.sh N': BASELINE configs 3, 4, 5 (and the headline config) on N GPUs of one box, launched
# the way the driver launches bench.py (torchrun, one rank per GPU).  Lines land in gpurun_out/r2_config*_nN.json.
N=$1; mkdir -p gpurun_out; PORT=29540
run() { name=$1; shift; PORT=$((PORT+1)); ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" ) > gpurun_out/$name.log 2>&1; echo "exit $?" >> gpurun_out/$name.log; grep '^{' gpurun_out/$name.log > gpurun_out/$name.json; tail -4 gpurun_out/$name.log | cut -c1-160; }
run r2_config3_layer11_n$N --config 3 --layer 11 --fast
if [ "$N" = "8" ]; then run r2_config3_layer23_n$N --config 3 --layer 23 --fast; fi
run r2_config4_layer11_n$N --config 4 --layer 11 --fast
if [ "$N" = "8" ]; then run r2_config4_layer39_n$N --config 4 --layer 39 --fast; fi
run r2_config5_n$N --config 5 --steps 10 --warmup 3
run r2_config2_n$N --fast
if [ "$N" = "2" ]; then timeout 600 python -m pytest tests/test_gpu_boundary.py -m gpu -q -p no:cacheprovider -k two_devices > gpurun_out/two_devices.log 2>&1; tail -2 gpurun_out/two_devices.log; fi
python tools/show_bench.py gpurun_out/r2_config3_layer11_n$N.json gpurun_out/r2_config4_layer11_n$N.json gpurun_out/r2_config2_n$N.json
