#!/bin/bash
# reference: examples/rec/hybrid_ncf.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/run_hetu.py
heturun=${workdir}/../../bin/heturun
${heturun} -s 1 -w ${1:-2} python ${mainpy} --comm Hybrid
