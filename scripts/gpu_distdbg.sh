#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29546"
O=gpurun_out/dist_dbg.jsonl; : > $O
for d in 0; do
B200_P2P_DEBUG=$d $TR scripts/dist_ops.py --dims 256 256 256 2>/dev/null | grep '^{' >> $O
B200_P2P_DEBUG=$d $TR scripts/dist_ops.py --dims 64 64 64 2>/dev/null | grep '^{' >> $O
done
cut -c 1-230 $O
