T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
echo "== trace"; TXF_TRACE=1 timeout 200 $T --master-port 29540 tools/exp/overlap2.py 2>&1 | grep -E "^world|profiled|^trace"
