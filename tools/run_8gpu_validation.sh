#!/usr/bin/env bash
# One 8-GPU box: multi-rank correctness, the 1->8 flagship bench with the stock arms, the collective
# roofline numbers, the ps/worker layouts and the chief-failure restart at full GPT-2-medium size.
# Every step has its own timeout; outputs land in gpurun_out/ (copied to profiles/ afterwards).
set -u
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/n8_topo.txt 2>&1
timeout 400 python -m pytest tests/test_multi_gpu.py -q -k 8gpu > $O/n8_pytest_collectives.log 2>&1; echo "collectives rc=$?"
timeout 300 python -m pytest tests/test_launcher_gpu.py -q -k "dedicated_parameter_servers" > $O/n8_pytest_ps.log 2>&1; echo "ps 2+6 rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29701 bench.py --gpus 8 > $O/n8_bench.json 2> $O/n8_bench.err; echo "bench8 rc=$?"
timeout 300 $TR --nproc-per-node 8 --master-port 29702 tools/collective_bench.py --sizes-mb 64,672 > $O/n8_collective.jsonl 2> $O/n8_collective.err; echo "coll rc=$?"
timeout 200 $TR --nproc-per-node 8 --master-port 29703 examples/widedeep-ps/widedeep_train.py --steps 40 --batch 16384 > $O/n8_widedeep_symmetric.log 2>&1; echo "wd sym rc=$?"
timeout 200 $TR --nproc-per-node 8 --master-port 29704 examples/widedeep-ps/widedeep_train.py --steps 40 --batch 16384 --ps_ranks 6,7 > $O/n8_widedeep_2ps_6workers.log 2>&1; echo "wd 2+6 rc=$?"
timeout 400 python tools/restart_probe.py > $O/n8_gpt2_restart.json 2> $O/n8_gpt2_restart.err; echo "restart rc=$?"
tail -2 $O/n8_pytest_collectives.log $O/n8_pytest_ps.log
grep -h "done" $O/n8_widedeep_symmetric.log $O/n8_widedeep_2ps_6workers.log
cut -c1-400 $O/n8_bench.json
cat $O/n8_gpt2_restart.json
