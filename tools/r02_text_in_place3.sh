#!/bin/bash
# One GPU: CSR batch arrays from the context's block cache; featuriser scratch in the lane's arena vs the stream-ordered pool.
mkdir -p gpurun_out
tag=${1:-r02z4}
timeout 300 python -m pytest tests/test_ingest_async_gpu.py tests/test_sparse_gpu.py tests/test_narrow_csr_gpu.py tests/test_parity_edges_gpu.py -m gpu -x -q --timeout 200 2>&1 | tail -2
B="python bench.py --steps 3 --warmup 3 --sub none --no-cpu-baseline --e2e-steps 48 --workload"
run() { name=$1; wl=$2; shift 2; env "$@" timeout 200 $B $wl > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; echo "== $name"; python tools/show_bench.py gpurun_out/${tag}_$name.json 2>/dev/null | head -1; grep "twtml ingest" gpurun_out/${tag}_$name.err | tail -2
python - <<PY
import json
l=[x for x in open("gpurun_out/${tag}_$name.json").read().splitlines() if x.startswith("{")][-1]
s=sorted(json.loads(l)["e2e"]["step_ms"]); print("median %.2f  p90 %.2f  max %.2f  sum %.1f" % (s[len(s)//2], s[int(len(s)*.9)], s[-1], sum(s)))
PY
}
run arena_trace tweets_1m TWTML_TRACE_INGEST=1
run arena1 tweets_1m A=1
run scratchpool1 tweets_1m TWTML_FEAT_POOL=1
run arena2 tweets_1m A=1
run csr csr_2m A=1
run wide tweets_1m_wide A=1
