#!/bin/bash
# reference: examples/nlp/bert (BASELINE config 5: DP x TP x PP)
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../train_hetu_bert_hybrid.py
heturun=${workdir}/../../../../bin/heturun
${heturun} -w ${1:-8} python ${mainpy} --size large --pp 2 --tp 2 --dp 2
