#!/bin/bash
# Developer A/B: builds gaussian-splatterer_b200/libgsb200_base.so from the sources of git revision $1 (default HEAD) so a
# probe can be run against both libraries in one gpurun call (GSB_LIB=... selects the library, see _lib.py).
set -e
REV=${1:-HEAD}
TMP=$(mktemp -d)
git archive $REV gaussian-splatterer_b200 include | tar -x -C $TMP
make -C $TMP/gaussian-splatterer_b200 -j8 > /dev/null
cp $TMP/gaussian-splatterer_b200/libgsb200.so gaussian-splatterer_b200/libgsb200_base.so
rm -rf $TMP
echo "built gaussian-splatterer_b200/libgsb200_base.so from $REV"
