#!/bin/bash
# tools/multi_gpu_sweep.sh -- bucket allreduce variants at N GPUs (development tool).
#   N=2 bash tools/multi_gpu_sweep.sh
N=${N:-2}
PORT=29600
mkdir -p gpurun_out
run() {
  PORT=$((PORT+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --bucket-mb $1 > gpurun_out/sweep_last.json 2> gpurun_out/sweep_last.err
  if [ -s gpurun_out/sweep_last.json ]; then
    python - "$TAG" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/sweep_last.json").read().strip().splitlines()[-1])
print(sys.argv[1], "n", d["n_gpus"], "bucket_mb", d["config"]["bucket_mb"], "buckets", d["config"]["buckets"],
      "ms/step", round(d["ms_per_step"], 3), "comm_ms", round(d["breakdown_ms"]["bucket_allreduce_path"], 3),
      "busbw", round(d["allreduce"]["busbw_gbs"], 1), "algbw", round(d["allreduce"]["algbw_gbs"], 1))
PY
  else
    echo "$TAG FAILED"; tail -5 gpurun_out/sweep_last.err
  fi
}
if [ -n "$BUCKETS" ]; then
  for mb in $BUCKETS; do TAG=${SWEEPTAG:-auto} run $mb; done
  exit 0
fi
if [ "$QUICK" = "3" ]; then
  TAG=auto run 256; TAG=auto run 1024; TAG=auto run 64
  APEX_B200_SYMM_CTAS=64 TAG=auto_ctas64 run 256
  exit 0
fi
if [ "$QUICK" = "2" ]; then
  APEX_B200_SYMM=0 TAG=nccl run 256
  APEX_B200_SYMM_P2P=0 TAG=nvls run 256
  APEX_B200_SYMM_P2P=1 TAG=p2p_ctas148 run 256
  TAG=auto run 256
  exit 0
fi
if [ "$QUICK" = "1" ]; then
  export APEX_B200_SYMM=0
  TAG=nccl run 256; TAG=nccl run 64
  export APEX_B200_SYMM=1
  TAG=nvls run 256
  APEX_B200_SYMM_CTAS=148 TAG=nvls_ctas148 run 256
  export APEX_B200_SYMM_P2P=1
  APEX_B200_SYMM_CTAS=148 TAG=p2p_ctas148 run 256
  APEX_B200_SYMM_CTAS=296 TAG=p2p_ctas296 run 256
  APEX_B200_SYMM_CTAS=296 TAG=p2p_ctas296 run 64
  exit 0
fi
export APEX_B200_SYMM=0
TAG=nccl run 256; TAG=nccl run 64; TAG=nccl run 20
export APEX_B200_SYMM=1
TAG=nvls run 256; TAG=nvls run 64; TAG=nvls run 20
for c in 16 32 128 296; do APEX_B200_SYMM_CTAS=$c TAG=nvls_ctas$c run 256; done
export APEX_B200_SYMM_P2P=1
TAG=p2p run 256
for c in 32 148 296; do APEX_B200_SYMM_CTAS=$c TAG=p2p_ctas$c run 256; done
