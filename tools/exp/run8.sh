T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 500 $T --master-port 29532 bench.py --gpus 8 --steps 20 --warmup 5 --e2e-steps 6 2>gpurun_out/b8.err | tail -1 > gpurun_out/r2_bench_8gpu.json
python -c "import json; d=json.load(open('gpurun_out/r2_bench_8gpu.json')); print(d['ms_per_step'], d['value'], d['parity_check']['ok'], d['e2e']['value'], d['e2e']['copy_gbs_per_gpu_each_way'])"; tail -2 gpurun_out/b8.err
echo "== trace"; TXF_TRACE=1 timeout 200 $T --master-port 29540 tools/exp/overlap2.py 2>&1 | grep -E "^world|profiled|^trace" | tee gpurun_out/r2_exchange_timeline_8gpu.txt
timeout 400 $T --master-port 29533 bench.py --gpus 8 --steps 10 --warmup 3 --no-e2e --global-rows 1000000000 2>gpurun_out/b8g.err | tail -1 > gpurun_out/r2_bench_8gpu_global1B.json
python -c "import json; d=json.load(open('gpurun_out/r2_bench_8gpu_global1B.json')); print(d['ms_per_step'], d['value'], d['config']['rows_per_gpu'], d['parity_check']['ok'])"; tail -2 gpurun_out/b8g.err
