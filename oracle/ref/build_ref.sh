#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Compiles the UNMODIFIED reference hot path (ctu-mrs/vns-sop) from the
# sources where they lie under $REF (default /root/reference) into oracle/_ref/:
#   oracle/_ref/sop_vns_ref        headless, seedable reference solver binary
#   oracle/_ref/liboracle_ref.so   extern "C" wrapper around op::VnsSopPath / op::VnsSopSolver
# No reference source is copied into the repo; the include tree is a set of symlinks in a temp dir.
# The reference's own build system (make + Boost + log4cxx + cairo + X11) is NOT run; the
# missing third-party headers are replaced by oracle/ref/shim/ (logging macros, BOOST_FOREACH,
# boost::mutex names).  Flags follow sources/Mk/flags.mk:3,11 (-O2).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/../_ref"
SRC="$REF/sources"
CRL="$SRC/comrob/crl"
if [ ! -d "$SRC" ]; then
   echo "build_ref: $SRC not present (GPU box?) - keeping prebuilt oracle/_ref" >&2
   exit 0
fi
mkdir -p "$OUT"
TMP="$(mktemp -d /tmp/oracle_ref.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
INC="$TMP/inc"
mkdir -p "$INC/crl"
for h in "$CRL"/crl/src/*.h; do ln -s "$h" "$INC/crl/$(basename "$h")"; done
ln -s "$CRL/crl.gui/src" "$INC/crl/gui"
ln -s "$CRL/crl.algorithm/src" "$INC/crl/alg"
ln -s "$CRL/crl.loader/src" "$INC/crl/loader"

CXX="${ORACLE_CXX:-/usr/bin/g++}"   # not $CXX: the image exports a g++ that links libstdc++ statically
FLAGS="-std=c++17 -O2 -w -fPIC -DSOP_VNS -I$HERE/shim -I$INC -I$SRC -I$HERE"
TUS=""
for f in vns_sop_path vns_sop_solver sop_loader dataset_loader_sop dataset_loader_op math_common \
         heuristic_types coords convexhull; do TUS="$TUS $SRC/$f.cpp"; done
for f in algorithm result_log text_result_log; do TUS="$TUS $CRL/crl.algorithm/src/$f.cc"; done
for f in timerN file_utils random bbox waitobject; do TUS="$TUS $CRL/crl/src/$f.cc"; done
for f in canvas shape shapes colormap colors xcolors view_transform renderer_object guihandler \
         guievent mouseclick waitforclick; do TUS="$TUS $CRL/crl.gui/src/$f.cc"; done

OBJS=""
pids=()
for tu in $TUS; do
   o="$TMP/$(basename "$tu" | tr . _).o"
   OBJS="$OBJS $o"
   $CXX $FLAGS -c "$tu" -o "$o" &
   pids+=($!)
done
# the wrapper includes vns_sop_solver.h with private→public to reach the neighbourhood drivers
$CXX $FLAGS -c "$HERE/ref_wrap.cpp" -o "$TMP/ref_wrap.o" &
pids+=($!)
$CXX $FLAGS -c "$HERE/ref_main.cpp" -o "$TMP/ref_main.o" &
pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done

$CXX -o "$OUT/sop_vns_ref" "$TMP/ref_main.o" $OBJS -lpthread
$CXX -shared -o "$OUT/liboracle_ref.so" "$TMP/ref_wrap.o" $OBJS -lpthread
echo "build_ref: wrote $OUT/sop_vns_ref and $OUT/liboracle_ref.so"
