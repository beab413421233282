#!/bin/bash
# Everything that is measured per GPU count, in one go (one gpurun call per N):
#   bash bench/run_scaling_suite.sh N [tests: all|core|none] [models: yes|no]
# Writes gpurun_out/suite_${N}gpu/*: multi-GPU test log, headline bench (both arms), collective sweep 1 KiB - 1 GiB vs NCCL with
# clock records, and the three north-star model configs (100 timed steps, clocks in every line, both arms).
N=${1:-2}; TESTS=${2:-all}; MODELS=${3:-yes}
OUT=gpurun_out/suite_${N}gpu
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "1" ]; then TR="python"; fi
if [ "$TESTS" = "all" ]; then
  timeout 600 python -m pytest tests -m "gpu and multigpu" -q 2>&1 | tail -12 > $OUT/tests_multigpu.log
elif [ "$TESTS" = "core" ]; then
  timeout 400 python -m pytest tests/test_p2p_collectives.py tests/test_convnet_trainer_ddp_gpu.py tests/test_syncbn_gpu.py tests/test_zero_gpu.py -q 2>&1 | tail -12 > $OUT/tests_multigpu.log
fi
timeout 200 $TR bench.py --gpus $N --steps 1000 --warmup 50 2> $OUT/bench_ours.err | tail -1 > $OUT/bench_ours.json
timeout 200 $TR bench.py --impl reference --gpus $N --steps 300 --warmup 20 2> $OUT/bench_ref.err | tail -1 > $OUT/bench_ref.json
if [ "$N" != "1" ]; then
  timeout 600 python bench/coll_sweep.py --gpus $N --max-mb 1024 --out $OUT/coll_sweep.jsonl > $OUT/coll_sweep.log 2>&1
fi
if [ "$MODELS" = "yes" ]; then
  for impl in ours torch; do
    timeout 300 python bench/train_models.py --model resnet18 --gpus $N --steps 100 --warmup 10 --impl $impl --out $OUT/models.jsonl > $OUT/m_r18_$impl.log 2>&1
    timeout 300 python bench/train_models.py --model resnet50 --bf16 --sync-bn --batch 64 --gpus $N --steps 100 --warmup 10 --impl $impl --out $OUT/models.jsonl > $OUT/m_r50_$impl.log 2>&1
    timeout 300 python bench/train_models.py --model gpt2 --bf16 --batch 8 --gpus $N --steps 100 --warmup 10 --impl $impl --out $OUT/models.jsonl > $OUT/m_gpt2_$impl.log 2>&1
  done
fi
cat $OUT/tests_multigpu.log 2>/dev/null | tail -3
cat $OUT/bench_ours.json | cut -c1-160; cat $OUT/bench_ref.json | cut -c1-160
wc -l $OUT/coll_sweep.jsonl 2>/dev/null; cat $OUT/models.jsonl 2>/dev/null | cut -c1-260
