#!/bin/bash
# One GPU: featuriser / ingest tests, then configs[0] end to end with the text column read in place vs copied.
mkdir -p gpurun_out
tag=${1:-r02t}
timeout 400 python -m pytest tests/test_ingest_async_gpu.py tests/test_sparse_gpu.py tests/test_parity_edges_gpu.py -m gpu -x -q --timeout 200 > gpurun_out/${tag}_pytest.log 2>&1; grep -n "^E " gpurun_out/${tag}_pytest.log | cut -c1-300 | head -30; tail -3 gpurun_out/${tag}_pytest.log
B="python bench.py --steps 5 --warmup 3 --sub none --no-cpu-baseline --e2e-steps 20"
timeout 200 $B --workload tweets_1m > gpurun_out/${tag}_tweets_in_place.json 2> gpurun_out/${tag}_a.err
TWTML_NO_TEXT_IN_PLACE=1 timeout 200 $B --workload tweets_1m > gpurun_out/${tag}_tweets_copied.json 2> gpurun_out/${tag}_b.err
timeout 200 $B --workload tweets_1m_wide > gpurun_out/${tag}_tweets_wide_in_place.json 2> gpurun_out/${tag}_c.err
for f in tweets_in_place tweets_copied tweets_wide_in_place; do echo $f; python tools/show_bench.py gpurun_out/${tag}_$f.json | head -1; python - <<PY
import json
l=[x for x in open("gpurun_out/${tag}_$f.json").read().splitlines() if x.startswith("{")][-1]
e=json.loads(l)["e2e"]; print({k:e[k] for k in ("h2d_bytes_per_step","ingest_ms_per_step","compute_ms_per_step","steps")})
PY
done
