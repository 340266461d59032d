#!/bin/bash
# quick GPU check of an extraction-kernel change: parity tests of the extraction path, then a short device-resident bench
# usage (on the GPU box): bash tools/gpu_quick.sh [tag]
tag=${1:-q}
timeout 600 python -m pytest tests/test_gpu_extract.py tests/test_gpu_dropin.py tests/test_gpu_orbslam2_mode.py -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/bq_$tag.json 2> gpurun_out/bq_$tag.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bq_$tag.json").read().strip().splitlines()[-1])
print("$tag", round(d["value"]), d["ms_per_step"], round(d["roofline"]["ms_per_step"], 4), d["kernel_ms_per_step"])
PY
