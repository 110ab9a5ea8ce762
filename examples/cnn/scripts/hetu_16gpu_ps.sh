#!/bin/bash
# reference: examples/cnn/scripts/hetu_16gpu_ps.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../main.py
heturun=${workdir}/../../../bin/heturun
${heturun} -c ${workdir}/hetu_config16ps.yml python ${mainpy} --model ${1:-resnet18} --dataset ${2:-CIFAR10} --learning-rate 0.01 --validate --timing --comm-mode PS
