#!/bin/bash
# reference: examples/cnn/scripts/hetu_8gpu.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../main.py
heturun=${workdir}/../../../bin/heturun
# 8 data-parallel workers, gradients all-reduced (lr pre-divided by 8 as in the reference)
${heturun} -w 8 python ${mainpy} --model ${1:-resnet18} --dataset ${2:-CIFAR10} --learning-rate 0.00125 --validate --timing --comm-mode AllReduce
