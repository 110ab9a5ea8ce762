#!/bin/bash
# reference: examples/cnn/scripts/hetu_16gpu.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../main.py
heturun=${workdir}/../../../bin/heturun
# two 8-GPU hosts described in hetu_config16allreduce.yml
${heturun} -c ${workdir}/hetu_config16allreduce.yml python ${mainpy} --model ${1:-resnet18} --dataset ${2:-CIFAR10} --learning-rate 0.000625 --validate --timing --comm-mode AllReduce
