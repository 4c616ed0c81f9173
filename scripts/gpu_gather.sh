#!/bin/bash
# gather of the packed slabs at N GPUs under a few NCCL settings (VARS="A=1;B=2 C=3"), N from $N
mkdir -p gpurun_out
N=${N:-4}
IFS=';' read -ra SETS <<< "${VARS:-;}"
for s in "${SETS[@]}"; do
  env $s timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 3 --warmup 2 --no-e2e > gpurun_out/gather.json 2> gpurun_out/gather.err || tail -3 gpurun_out/gather.err
  python - "$s" <<'PY'
import json,sys
try:
    d=json.loads(open("gpurun_out/gather.json").read().strip().splitlines()[-1])
    print(sys.argv[1] or "(default)", "| value %.0f | gather" % d["value"], {k: d["gather"][k] for k in ("ms","gbs","frac_of_900_gbs_ingress","sink_slab_ok")})
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
