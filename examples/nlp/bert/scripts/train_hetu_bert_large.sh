#!/bin/bash
# reference: examples/nlp/bert/scripts/train_hetu_bert_large.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../train_hetu_bert.py
heturun=${workdir}/../../../../bin/heturun
python ${mainpy} --size large --batch-size 32 --seq-len 128 --lr 1e-5 --steps ${1:-100}
