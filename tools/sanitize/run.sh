#!/bin/bash
# Host code of the library under AddressSanitizer + UndefinedBehaviorSanitizer
# (+ leak check): the visibility ordering on 300 random box tables (cycle
# breaking and DOT dump included, plus footprints, overlay edges and the automatic camera on random cameras) and the plotfile reader on 400 corrupted
# plotfiles, then 2000 random convexify calls.  CPU only; ~2 minutes.  Last run (round 1): no findings.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
src="$here/../../amrvolumerenderer_b200/csrc"
work="${1:-/tmp/amrvr_sanitize}"
mkdir -p "$work"
FLAGS="-std=c++17 -O1 -g -fsanitize=address,undefined -fno-omit-frame-pointer -mavx2 -I$src"
g++ $FLAGS "$here/order_harness.cpp" "$src/host_logic.cpp" -o "$work/order_harness" -lpthread
g++ $FLAGS "$here/plotfile_harness.cpp" "$src/amr_io.cpp" -o "$work/plotfile_harness" -lpthread
python "$here/gen_corpus.py" "$work/corpus"
"$work/order_harness"
"$work/plotfile_harness" "$work/corpus/base" "$work"/corpus/m*
