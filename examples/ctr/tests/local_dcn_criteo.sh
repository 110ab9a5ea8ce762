#!/bin/bash
# reference: examples/ctr/tests/local_dcn_criteo.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../run_hetu.py
heturun=${workdir}/../../../bin/heturun
python ${mainpy} --model dcn_criteo --nepoch ${1:-1} --all
