for v in ${VARIANTS:-default}; do
  if [ $v = default ]; then unset SFMBA_LIB; else export SFMBA_LIB=$PWD/variants/libsfmba_$v.so; fi
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-extras 2>/dev/null | python -c "
import sys, json
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', round(j['value'],1), {k: round(v,4) for k,v in j['phase_ms_per_step'].items()}, round(j['kernel_ms']['pcg_tail'],4))"
done
