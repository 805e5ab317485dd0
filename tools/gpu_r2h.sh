#!/bin/bash
# round 2, GPU call H (8 GPUs): bench at N = 8 and N = 4 (default NCCL settings, and with more channels for the send / recv all-gather)
set -x
mkdir -p gpurun_out
run() {  # name, nproc, extra env
  local name=$1 n=$2; shift 2
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/$name.json 2> gpurun_out/$name.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms', round(d['ms_per_step'],3), 'value', round(d['value']), 'ag_ms', round(d.get('allgather_B_ms',0),3), 'inbound GB/s', round(d.get('allgather_inbound_gbs',0)), {k:round(v,3) for k,v in d.items() if k.startswith('allgather_plus')}, 'with_ag', round(d.get('value_with_allgather',0)), 'e2e ms', round(d['e2e']['ms_per_step'],2), d.get('sharded_entry_matches_replicated'))
" || tail -5 gpurun_out/$name.err
}
run r2h_bench_n8 8 NCCL_DEBUG=WARN
run r2h_bench_n8_ch32 8 NCCL_DEBUG=WARN NCCL_MIN_NCHANNELS=32
run r2h_bench_n4 4 NCCL_DEBUG=WARN
