#!/bin/bash
# Variants of the data-parallel gradient exchange (tools/overlap_probe.py), one torchrun each; N = $1 ranks.
# usage: tools/overlap_sweep.sh N [full]
N=${1:-2}
OUT=gpurun_out/overlap_sweep_n$N.jsonl
mkdir -p gpurun_out
: > $OUT
run() {
  label=$1; shift
  envs=(); args=()
  for a in "$@"; do if [[ $a == *=* && $a != --* ]]; then envs+=("$a"); else args+=("$a"); fi; done
  env "${envs[@]}" timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port 29541 tools/overlap_probe.py --label $label "${args[@]}" 2>gpurun_out/overlap_$label.err | grep '^{' >> $OUT
  echo "$label rc=$?"
}
if [ "$2" != "only-full" ] && [ "$2" != "n8" ]; then
run none --mode none
run after --mode after
run buckets --timeline gpurun_out/timeline_buckets_n$N.json
run buckets_b32 --bucket-mb 32
run buckets_b16 --bucket-mb 16
fi
if [ "$2" == "full" ] || [ "$2" == "only-full" ]; then
  run b32_ctas16 NCCL_MAX_CTAS=16 --bucket-mb 32
  run b32_ctas16_sm132 NCCL_MAX_CTAS=16 B200SD_SM_LIMIT=132 --bucket-mb 32
  run b32_ctas8_sm140 NCCL_MAX_CTAS=8 B200SD_SM_LIMIT=140 --bucket-mb 32
  run b32_thr256 NCCL_NTHREADS=256 --bucket-mb 32
fi
if [ "$2" == "n8" ]; then
  run none --mode none
  run after --mode after
  run b32 --bucket-mb 32 --timeline gpurun_out/timeline_b32_n$N.json
  run b32_thr256 NCCL_NTHREADS=256 --bucket-mb 32 --timeline gpurun_out/timeline_b32_thr256_n$N.json
  run b32_ctas8_sm140 NCCL_MAX_CTAS=8 B200SD_SM_LIMIT=140 --bucket-mb 32
  run b32_thr256_ctas16 NCCL_NTHREADS=256 NCCL_MAX_CTAS=16 --bucket-mb 32
fi
cat $OUT | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); t = d.get('timeline', {})
    print(d['label'], d['ms_per_step'], d.get('env'), 'buckets', d.get('buckets'), {k: v for k, v in t.items() if k != 'nccl_intervals_us'})
    print('   nccl at', t.get('nccl_intervals_us'))
"
