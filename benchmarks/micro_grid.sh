#!/bin/bash
# BASELINE.json config 5: codec x size x world-size grid, ours (native plans) next to the unmodified reference
# (grace_dl.dist over NCCL, baseline/_ref) and plain NCCL all-reduce, every row with its own nvidia-smi clock record.
#   bash benchmarks/micro_grid.sh <world> [out_dir] [sizes...]
# Writes <out_dir>/micro_grid_w<world>_<size>_{ours,reference}.json; benchmarks/micro_grid_table.py merges them.
set -u
W="${1:-1}"; OUT="${2:-gpurun_out}"; shift 2 2>/dev/null || true
SIZES="${*:-1048576 16777216 134217728 1073741824}"
HERE="$(cd "$(dirname "$0")" && pwd)"
mkdir -p "$OUT"
run() {  # run <script> <args...>
  if [ "$W" -gt 1 ]; then
    python -m torch.distributed.run --nnodes=1 --nproc-per-node "$W" --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) "$@"
  else
    python "$@"
  fi
}
for N in $SIZES; do
  IT=10; RIT=3; RCODECS="none,topk,qsgd,signsgd,powersgd"
  if [ "$N" -ge 1000000000 ]; then IT=4; RIT=2; RCODECS="none,topk,qsgd"; fi
  echo "=== $N elements, W=$W: ours"
  GRACE_B200_ARENA_MB=2048 run "$HERE/micro.py" --shape "flat:$N" --iters $IT --warmup 2 \
      --codecs none,fp16,topk,randomk,threshold,dgc,qsgd,terngrad,natural,u8bit,signsgd,signum,efsignsgd,onebit,powersgd,nvls,nccl \
      --out "$OUT/micro_grid_w${W}_${N}_ours.json" 2>&1 | grep -E "W=$W|skipped|status|rror"
  echo "=== $N elements, W=$W: reference"
  run "$HERE/micro_reference.py" --shape "flat:$N" --iters $RIT --warmup 1 --codecs "$RCODECS" \
      --out "$OUT/micro_grid_w${W}_${N}_reference.json" 2>&1 | grep -E "W=$W|fails|rror"
done
