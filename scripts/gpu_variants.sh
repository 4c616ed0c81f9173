#!/bin/bash
# quick kernel-only bench of every library variant under build/ (MNWCUDA_LIB override), then the in-tree one
mkdir -p gpurun_out
for lib in "" build/libminnow_cuda_*.so; do
  tag=$(basename "${lib:-intree}" .so)
  MNWCUDA_LIB=${lib:+$PWD/$lib} timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu --no-e2e --no-sweep $QARGS > gpurun_out/var_$tag.json 2> gpurun_out/var_$tag.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/var_%s.json" % tag).read().strip().splitlines()[-1])
    print("%-28s enc %.3f ms dec %.3f ms k_encode %.3f ok=%s" % (tag, d["encode_ms"], d["decode_ms"], d["kernel_ms"]["k_encode"], d["check"]["ok"]))
except Exception as e:
    print(tag, "unreadable", e); print(open("gpurun_out/var_%s.err" % tag).read()[-1500:])
PY
done
