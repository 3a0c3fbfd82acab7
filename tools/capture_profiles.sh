#!/usr/bin/env bash
# Reproduce the evidence under profiles/ on one B200 box (run from the repo root; every ncu step follows a plain run
# of the same command, and no number printed under ncu is used as a benchmark value).
#   bash tools/capture_profiles.sh [out_dir]          # default gpurun_out/
set -u
O=${1:-gpurun_out}
mkdir -p "$O"
# 1. headline: ours + both stock arms + ResNet-50 + launcher latency in one JSON line
python bench.py > "$O/bench_default_n1.json" 2> "$O/bench_default_n1.err"
# 2. kernel-level comparisons
python tools/microbench.py > "$O/microbench.jsonl" 2> "$O/microbench.err"
python tools/gemm_bench.py > "$O/gemm_vs_cublas_per_shape.jsonl" 2>&1
python tools/gemm_bench.py resnet > "$O/gemm_vs_cublas_resnet_pointwise.jsonl" 2>&1
python tools/attn_bench.py > "$O/attention_vs_library.jsonl" 2>&1
python tools/conv_bench.py --md "$O/conv_per_layer_vs_cudnn.md" > "$O/conv_per_layer_vs_cudnn.jsonl" 2> "$O/conv.err"
# 3. ncu --set full of the top kernels (plain run first)
python tools/ncu_pack.py > "$O/plain.log" 2>&1 && python tools/conv_probe.py >> "$O/plain.log" 2>&1 &&
    python tools/gemm_probe.py >> "$O/plain.log" 2>&1 || { echo "plain probe run failed"; exit 1; }
NCU="ncu --set full --clock-control none --import-source on"
$NCU -k regex:"layernorm|softmax_xent|colsum|flat_optim|fused_reduce_optim" -c 40 -o "$O/pack" python tools/ncu_pack.py > "$O/ncu1.log" 2>&1
$NCU -k regex:"conv_bf16|conv3x3" -c 8 -o "$O/conv" python tools/conv_probe.py > "$O/ncu2.log" 2>&1
$NCU -k regex:"gemm_bf16|splitk" -c 8 -o "$O/gemm" python tools/gemm_probe.py > "$O/ncu3.log" 2>&1
# 4. per-launch durations of whole steps, warm caches
L="ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv"
$L -c 9000 --log-file "$O/launches_bert_large_n1_warm.csv" python bench.py --steps 1 --warmup 3 --compare 0 --secondary 0 --skip-e2e > "$O/ncu4.log" 2>&1
$L -c 4000 --log-file "$O/launches_resnet50_n1_warm.csv" python bench.py --model resnet50 --steps 1 --warmup 3 --compare 0 --skip-e2e > "$O/ncu5.log" 2>&1
# 5. summaries (these also run without a GPU)
python tools/ncu_summary.py "$O/pack.ncu-rep" --max 14 --out "$O/ncu_memory_bound_kernels.md"
python tools/ncu_summary.py "$O/conv.ncu-rep" --max 8 --out "$O/ncu_conv.md"
python tools/ncu_summary.py "$O/gemm.ncu-rep" --pick 1,3,6,7 --out "$O/ncu_gemm.md"
python tools/launch_summary.py "$O/launches_bert_large_n1_warm.csv" > "$O/launches_bert_large_n1_warm_summary.txt"
python tools/launch_summary.py "$O/launches_resnet50_n1_warm.csv" > "$O/launches_resnet50_n1_warm_summary.txt"
python tools/sass_table.py --out "$O/sass_opcode_table.md"
echo "done: see $O/"
