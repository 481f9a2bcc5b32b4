#!/usr/bin/env bash
# A/B of the chunk-stream k_freq_gemm against the barrier-synchronised tile-by-tile form it replaces (v18), on the tricylinder
# geometry at several mode counts and at the full big_tricylinder size.  Run on a GPU box:  bash tools/experiments/ab_stream.sh
# stream_check.py prints the stage times of the default plan ("variant") and of a plan created under the given environment
# switches ("baseline"), and the largest relative difference of every output between the two.
ALLOFF="MEEMB200_NO_STREAM=1 MEEMB200_NO_STREAM_SYRK=1 MEEMB200_NO_BATCH_INV=1"
for nm in 72 75 100 104 128 150 200 224 248 256; do
    echo "== $nm modes per region, 512 frequencies"
    timeout 120 python tools/experiments/stream_check.py 512 $nm $ALLOFF 2>&1 | tail -7 | cut -c1-220
done
echo "== big_tricylinder (200 modes, 4096 frequencies)"
timeout 200 python tools/experiments/stream_check.py 4096 200 $ALLOFF 2>&1 | tail -7
