#!/bin/bash
# One GPU: where the tweet ingest spends its time beside a training step (TWTML_TRACE_INGEST), with and without stream
# priorities, long- and short-lived featuriser CTAs, text in place or copied.
mkdir -p gpurun_out
tag=${1:-r02u}
B="python bench.py --steps 3 --warmup 3 --sub none --no-cpu-baseline --e2e-steps 12 --workload tweets_1m"
run() { name=$1; shift; env TWTML_TRACE_INGEST=1 "$@" timeout 200 $B > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; echo "== $name"; python tools/show_bench.py gpurun_out/${tag}_$name.json | head -1; grep "twtml ingest" gpurun_out/${tag}_$name.err | tail -4; }
run noprio TWTML_NO_STREAM_PRIORITY=1
run prio
run prio_short TWTML_FEAT_SHORT=1
run prio_copied TWTML_NO_TEXT_IN_PLACE=1
