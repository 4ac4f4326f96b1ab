#!/bin/bash
# round 2, GPU call K (2 GPUs): where does a distributed pattern-kernel launch lose its time?  level-0 operators at
# 2 ranks vs the same rows on 1 rank, with the protocol steps switched off one by one (timing only, results invalid)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29514"
timeout 300 python scripts/dist_ops.py --dims 256 256 128 --reps 50 > gpurun_out/r2k_ops_1rank_halfrows.json 2> gpurun_out/r2k_ops_1rank.err; cat gpurun_out/r2k_ops_1rank_halfrows.json
timeout 300 $TR --nproc-per-node 2 scripts/dist_ops.py --dims 256 256 256 --reps 50 > gpurun_out/r2k_ops_2rank.json 2> gpurun_out/r2k_ops_2rank.err; grep "^{" gpurun_out/r2k_ops_2rank.json
for dbg in 2 6 14; do
B200_P2P_DEBUG=$dbg timeout 300 $TR --nproc-per-node 2 scripts/dist_ops.py --dims 256 256 256 --reps 50 > gpurun_out/r2k_ops_2rank_dbg$dbg.json 2> gpurun_out/r2k_ops_2rank_dbg$dbg.err; echo "dbg=$dbg"; grep "^{" gpurun_out/r2k_ops_2rank_dbg$dbg.json
done
B200_CHUNKED=0 timeout 300 $TR --nproc-per-node 2 scripts/dist_ops.py --dims 256 256 256 --reps 50 > gpurun_out/r2k_ops_2rank_strided.json 2> gpurun_out/r2k_ops_2rank_strided.err; echo strided; grep "^{" gpurun_out/r2k_ops_2rank_strided.json
B200_COMPRESS=1 timeout 300 $TR --nproc-per-node 2 scripts/dist_ops.py --dims 256 256 256 --reps 50 > gpurun_out/r2k_ops_2rank_val8.json 2> gpurun_out/r2k_ops_2rank_val8.err; echo val8; grep "^{" gpurun_out/r2k_ops_2rank_val8.json
B200_HALO=nccl timeout 300 $TR --nproc-per-node 2 scripts/dist_ops.py --dims 256 256 256 --reps 50 > gpurun_out/r2k_ops_2rank_nccl.json 2> gpurun_out/r2k_ops_2rank_nccl.err; echo nccl; grep "^{" gpurun_out/r2k_ops_2rank_nccl.json
