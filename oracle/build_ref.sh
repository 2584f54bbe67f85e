#!/bin/bash
# TEST INFRASTRUCTURE ONLY.  Builds the reference's own sw runtime (python_sw) for the
# five in-scope networks from the UNMODIFIED sources where they lie under
# /root/reference, with the make-sw.sh:114 flags, against oracle/shim stand-in headers.
# Also builds oracle/_ref/tap-<net>.so, the per-layer tap harness (oracle/tap.cpp) over the same unmodified sources.
# Outputs only into oracle/_ref/ (git-ignored).  Usage: oracle/build_ref.sh [net ...]
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=${BNN_REFERENCE:-/root/reference}
SRC=$REF/bnn/src
OUT=$HERE/_ref
mkdir -p "$OUT"
NETS=${@:-cnvW1A1 cnvW1A2 cnvW2A2 lfcW1A1 lfcW1A2}
OPT=${REF_OPT:--O2}
for n in $NETS; do
  ( g++ -g0 -DOFFLOAD -DRAWHLS -std=c++14 -pthread $OPT -fPIC -shared -w \
      $SRC/library/host/foldedmv-offload.cpp $SRC/library/host/rawhls-offload.cpp \
      $SRC/network/$n/hw/top.cpp $SRC/network/$n/sw/main_python.cpp $HERE/shim/bbj_forward.cpp $HERE/shim/ref_hooks.cpp \
      -I$HERE/shim -I$SRC/library/host -I$SRC/library/finn-hlslib -I$SRC/network/$n/hw \
      -o $OUT/python_sw-$n.so && echo "built $OUT/python_sw-$n.so" ) &
  # per-layer tap harness: the same build with oracle/tap.cpp (which #includes the network's unmodified top.cpp) in its place
  ( g++ -g0 -DOFFLOAD -DRAWHLS -DTAP_NET_$n -std=c++14 -pthread $OPT -fPIC -shared -w \
      $SRC/library/host/foldedmv-offload.cpp $SRC/library/host/rawhls-offload.cpp \
      $HERE/tap.cpp $SRC/network/$n/sw/main_python.cpp $HERE/shim/bbj_forward.cpp \
      -I$HERE/shim -I$SRC/library/host -I$SRC/library/finn-hlslib -I$SRC/network/$n/hw \
      -o $OUT/tap-$n.so && echo "built $OUT/tap-$n.so" ) &
done
wait
