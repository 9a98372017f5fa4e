# gpurun --gpus 8 -- 'bash tools/run_scale8.sh': the driver's N=8 launch of both arms
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_n8.log 2> gpurun_out/bench_n8.err; echo "exit $?" >> gpurun_out/bench_n8.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/bench_ref_n8.log 2> gpurun_out/bench_ref_n8.err; echo "exit $?" >> gpurun_out/bench_ref_n8.log
tail -c 600 gpurun_out/bench_n8.log; tail -c 400 gpurun_out/bench_ref_n8.log
