#!/bin/bash
# reference: examples/moe/scripts/run_top1_16gpus.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../test_moe.py
heturun=${workdir}/../../../bin/heturun
${heturun} -c ${workdir}/../../cnn/scripts/hetu_config16allreduce.yml python ${mainpy} --gate top --top 1 --batch-size 16 --num-tokens 1024 --model-dim 2048 --hidden-size 2048 --num-local-experts 2
