set -x
cd /root/repo
nvidia-smi topo -m > gpurun_out/c51_topo.txt 2>&1
python -m pytest tests/test_multi_gpu.py -x -q -m gpu 2>&1 | tail -3
for N in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N --steps 8 --warmup 3 > gpurun_out/c51_bench_ours_n$N.json 2> gpurun_out/c51_bench_ours_n$N.err; tail -c 600 gpurun_out/c51_bench_ours_n$N.json
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29600 bench.py --impl stock --gpus 8 --steps 8 --warmup 3 > gpurun_out/c51_bench_stock_n8.json 2> gpurun_out/c51_bench_stock_n8.err; tail -c 300 gpurun_out/c51_bench_stock_n8.json
TONY_B200_SYMM=ipc python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/c51_bench_ours_n8_ipc.json 2> gpurun_out/c51_bench_ours_n8_ipc.err; tail -c 300 gpurun_out/c51_bench_ours_n8_ipc.json
