#!/bin/bash
# A/B of builds / environment switches under torchrun on one box:
#   NPROC=4 VARIANTS="default default:SFMBA_STEP1=0 fence1" bash profiles/ab_2gpu_variants.sh
# (a name is variants/libsfmba_<name>.so, `default` the in-tree build; `:VAR=VALUE` sets an environment switch)
N=${NPROC:-2}
for spec in ${VARIANTS:-default}; do
  v=${spec%%:*}; kv=""; [ "$spec" != "$v" ] && kv=${spec#*:}
  if [ $v = default ]; then unset SFMBA_LIB; else export SFMBA_LIB=$PWD/variants/libsfmba_$v.so; fi
  env $kv python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --no-extras 2>/dev/null | python -c "
import sys, json
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$spec', round(j['value'],1), {k: round(v,4) for k,v in j['phase_ms_per_step'].items()}, round(j['kernel_ms']['pcg_tail'],4))"
done
