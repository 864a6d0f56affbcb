T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
echo "== default"; timeout 200 $T --master-port 29540 tools/exp/overlap2.py 2>&1 | grep -E "^world|profiled"
echo "== no gather wait, 140 persistent CTAs"; TXF_NO_GATHER_WAIT=1 TXF_PERSISTENT_CTAS=140 timeout 200 $T --master-port 29541 tools/exp/overlap2.py 2>&1 | grep -E "^world|profiled"
echo "== gather wait, 140 persistent CTAs"; TXF_PERSISTENT_CTAS=140 timeout 200 $T --master-port 29542 tools/exp/overlap2.py 2>&1 | grep -E "^world|profiled"
