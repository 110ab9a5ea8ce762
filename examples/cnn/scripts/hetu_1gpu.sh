#!/bin/bash
# reference: examples/cnn/scripts/hetu_1gpu.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../main.py
heturun=${workdir}/../../../bin/heturun
# bash hetu_1gpu.sh mlp CIFAR10
python ${mainpy} --model ${1:-mlp} --dataset ${2:-CIFAR10} --learning-rate 0.01 --validate --timing
