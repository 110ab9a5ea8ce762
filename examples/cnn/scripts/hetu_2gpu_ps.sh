#!/bin/bash
# reference: examples/cnn/scripts/hetu_2gpu_ps.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../main.py
heturun=${workdir}/../../../bin/heturun
# 1 server + 2 workers through the parameter server
${heturun} -s 1 -w 2 python ${mainpy} --model ${1:-mlp} --dataset ${2:-CIFAR10} --learning-rate 0.01 --validate --timing --comm-mode PS
