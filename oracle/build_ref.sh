#!/usr/bin/env bash
# build_ref.sh -- compile the UNMODIFIED reference CPU path into oracle/_ref/.
# TEST INFRASTRUCTURE.  Sources are compiled where they lie under $REF (default
# /root/reference); nothing is copied.  The only substitutions are the MPI/PAPI
# stand-ins (oracle/stubs) and the force-included knob header (ref_config.h),
# because the reference's knobs are compile-time literals (core/Configuration.h:28-37).
# Flags follow the reference's own CMakeLists.txt:10,15 (-O3 -std=c++11 -mavx -DETH ...).
#
#   usage: build_ref.sh NET_BITS LOC_BITS PAYLOAD_BITS TWO_LEVEL
#   output: oracle/_ref/hpcjoin_ref_n<NET>_l<LOC>_p<PAYLOAD>_t<TWO_LEVEL>
set -euo pipefail
here="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
nb="${1:-5}"; lb="${2:-5}"; pb="${3:-27}"; tl="${4:-0}"
out="$here/_ref/hpcjoin_ref_n${nb}_l${lb}_p${pb}_t${tl}"
if [ ! -d "$REF" ]; then
  echo "reference tree $REF not present; keeping prebuilt $out" >&2
  [ -x "$out" ] && exit 0 || exit 1
fi
mkdir -p "$here/_ref"
srcs=(
  "$REF/operators/HashJoin.cpp"
  "$REF/data/Relation.cpp" "$REF/data/Window.cpp"
  "$REF/memory/Pool.cpp"
  "$REF/histograms/LocalHistogram.cpp" "$REF/histograms/GlobalHistogram.cpp"
  "$REF/histograms/AssignmentMap.cpp" "$REF/histograms/OffsetMap.cpp"
  "$REF/tasks/HistogramComputation.cpp" "$REF/tasks/NetworkPartitioning.cpp"
  "$REF/tasks/LocalPartitioning.cpp" "$REF/tasks/BuildProbe.cpp"
  "$REF/performance/Measurements.cpp"
)
g++ -O3 -std=c++11 -mavx -w -DETH \
  -DMEASUREMENT_DETAILS_HISTOGRAM -DMEASUREMENT_DETAILS_NETWORK \
  -DMEASUREMENT_DETAILS_LOCALPART -DMEASUREMENT_DETAILS_LOCALBP \
  -DREF_NET_BITS="$nb" -DREF_LOC_BITS="$lb" -DREF_PAYLOAD_BITS="$pb" -DREF_TWO_LEVEL="$tl" \
  -include "$here/ref_config.h" -I"$here/stubs" -I"$REF" \
  "${srcs[@]}" "$here/ref_harness.cpp" "$here/stubs/mpi_shim.cpp" \
  -lpthread -o "$out"
echo "$out"
