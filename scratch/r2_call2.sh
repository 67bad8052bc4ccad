#!/bin/bash
# Round 2, GPU call 2 (TWO B200):  gpurun --gpus 2 --timeout 1500 -- 'bash scratch/r2_call2.sh'
# Peer transport (NVLink peer stores + mailbox counters, no NCCL in the halo path) against the NCCL transport:
# parity on 2 slabs, then the bench line both ways.
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29533 tests/mgpu_worker.py > $O/mgpu_p2p.log 2>&1; echo "mgpu p2p rc=$?" | tee -a $O/mgpu_p2p.log
CZB_NO_P2P=1 timeout 600 $TR --master-port 29534 tests/mgpu_worker.py > $O/mgpu_nccl.log 2>&1; echo "mgpu nccl rc=$?" | tee -a $O/mgpu_nccl.log
for rep in 1 2 3; do
  timeout 600 $TR --master-port 2954$rep bench.py --gpus 2 --steps 20 --warmup 3 --no-ttc >> $O/bench2_p2p.jsonl 2>> $O/bench2.err; echo "bench p2p rc=$?"
done
CZB_NO_P2P=1 timeout 600 $TR --master-port 29551 bench.py --gpus 2 --steps 20 --warmup 3 --no-ttc >> $O/bench2_nccl.jsonl 2>> $O/bench2.err; echo "bench nccl rc=$?"
# 8-GPU-sized slabs on 2 ranks: 1024 x 256 x 1024 cut in two = the per-rank work of an 8-GPU run
CZB_BENCH_NY=256 timeout 600 $TR --master-port 29552 bench.py --gpus 2 --steps 20 --warmup 3 --no-ttc --no-also >> $O/bench2_ny256_p2p.jsonl 2>> $O/bench2.err
CZB_BENCH_NY=256 CZB_NO_P2P=1 timeout 600 $TR --master-port 29553 bench.py --gpus 2 --steps 20 --warmup 3 --no-ttc --no-also >> $O/bench2_ny256_nccl.jsonl 2>> $O/bench2.err
timeout 600 $TR --master-port 29554 bench.py --gpus 2 --steps 20 --warmup 3 --weak --no-ttc --no-also >> $O/bench2_weak.jsonl 2>> $O/bench2.err; echo "weak rc=$?"
grep -h "mgpu\|rc=" $O/mgpu_p2p.log $O/mgpu_nccl.log | tail -40
for f in bench2_p2p bench2_nccl bench2_ny256_p2p bench2_ny256_nccl bench2_weak; do echo "== $f"; cut -c1-420 $O/$f.jsonl; done
tail -5 $O/bench2.err
