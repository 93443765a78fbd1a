#!/bin/sh
# ORACLE — test infrastructure only.  Regenerates tests/golden/rng_kat.json from
# the reference's own lib/random.cuh (needs /root/reference; run in the build container).
set -e
cd "$(dirname "$0")"
make ref
./_ref/ref_rng > ../tests/golden/rng_kat.json
echo "wrote tests/golden/rng_kat.json"
