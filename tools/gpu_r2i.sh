#!/bin/bash
# round 2, GPU call I (N GPUs): peer-push all-gather vs NCCL send/recv.  usage: bash tools/gpu_r2i.sh N
N=${1:-2}
set -x
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
  timeout 600 python -m pytest tests/test_shard.py -x -q -m gpu > gpurun_out/r2i_pytest_shard_n2.log 2>&1
  tail -5 gpurun_out/r2i_pytest_shard_n2.log
fi
run() {
  local name=$1 n=$2; shift 2
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/$name.json 2> gpurun_out/$name.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms', round(d['ms_per_step'],3), 'value', round(d['value']), 'ag_ms', round(d.get('allgather_B_ms',0),3), 'inbound GB/s', round(d.get('allgather_inbound_gbs',0)), {k:round(v,3) for k,v in d.items() if k.startswith('allgather_plus')}, 'with_ag', round(d.get('value_with_allgather',0)), 'e2e ms', round(d['e2e']['ms_per_step'],2), d.get('sharded_entry_matches_replicated'), d.get('allgather_path'))
" || tail -5 gpurun_out/$name.err
}
run r2i_bench_n${N}_push $N NCCL_DEBUG=WARN SPMMDB_MG_PEER=2
run r2i_bench_n${N}_push_ctas592 $N NCCL_DEBUG=WARN SPMMDB_MG_PEER=2 SPMMDB_MG_PUSH_CTAS=592
