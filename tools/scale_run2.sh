set -x
cd /root/repo
run() { # N impl model tag
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29700+$1)) bench.py --impl $2 --model $3 --gpus $1 --steps 8 --warmup 4 > gpurun_out/c96_$4.json 2> gpurun_out/c96_$4.err; tail -c 400 gpurun_out/c96_$4.json
}
run 8 ours resnet50 resnet50_ours_n8
run 8 stock resnet50 resnet50_stock_n8
run 4 ours resnet50 resnet50_ours_n4
run 2 ours resnet50 resnet50_ours_n2
run 8 ours widedeep widedeep_ours_n8
