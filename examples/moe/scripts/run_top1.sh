#!/bin/bash
# reference: examples/moe/scripts/run_top1.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../test_moe.py
heturun=${workdir}/../../../bin/heturun
# 8 GPUs x 2 local experts, 16 x 1024 tokens per GPU, model_dim = hidden = 2048 (reference run_top1.sh)
NCCL_DEBUG=${NCCL_DEBUG:-WARN} ${heturun} -w ${1:-8} python ${mainpy} --gate top --top 1 --batch-size 16 --num-tokens 1024 --model-dim 2048 --hidden-size 2048 --num-local-experts 2
