#!/bin/bash
# reference: examples/nlp/bert/scripts/train_hetu_bert_base_dp.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../train_hetu_bert.py
heturun=${workdir}/../../../../bin/heturun
# data parallel over N GPUs, batch 32 per GPU, Adam lr 1e-5 (reference train_hetu_bert_base_dp.sh)
${heturun} -w ${1:-8} python ${mainpy} --size base --batch-size 32 --seq-len 128 --lr 1e-5 --steps ${2:-100} --dp
