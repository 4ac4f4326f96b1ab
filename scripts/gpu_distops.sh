#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29546"
O=gpurun_out/dist_ops.jsonl; : > $O
python scripts/dist_ops.py --dims 256 256 128 2>/dev/null | tail -n 1 >> $O
$TR scripts/dist_ops.py --dims 256 256 256 2>/dev/null | grep '^{' >> $O
B200_HALO=nccl $TR scripts/dist_ops.py --dims 256 256 256 2>/dev/null | grep '^{' >> $O
python scripts/dist_ops.py --dims 100 100 50 2>/dev/null | tail -n 1 >> $O
$TR scripts/dist_ops.py --dims 100 100 100 2>/dev/null | grep '^{' >> $O
B200_HALO=nccl $TR scripts/dist_ops.py --dims 100 100 100 2>/dev/null | grep '^{' >> $O
python scripts/dist_ops.py --dims 32 32 16 2>/dev/null | tail -n 1 >> $O
$TR scripts/dist_ops.py --dims 32 32 32 2>/dev/null | grep '^{' >> $O
cat $O
