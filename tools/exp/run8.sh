N=$1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 500 $T --master-port 29532 bench.py --gpus $N --steps 20 --warmup 5 --e2e-steps 6 2>gpurun_out/b$N.err | tail -1 > gpurun_out/r2_bench_${N}gpu.json
python -c "import json; d=json.load(open('gpurun_out/r2_bench_${N}gpu.json')); print(d['ms_per_step'], d['value'], d['parity_check']['ok'], d['phases_ms'], d['e2e']['value'], d['e2e']['frac_of_ceiling'], d['e2e']['pcie_ceiling'])"; tail -2 gpurun_out/b$N.err
