#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY (see oracle/README.md).
# Compiles the reference's own C extension `_mfv2d` straight from the sources where they lie
# under /root/reference with plain gcc (no cmake, nothing copied into this repository).
# Output: oracle/_ref/_mfv2d_ref.so  (git-ignored; it still travels to the GPU box).
# Source list = MFV2D_SOURCE_FILES of the reference's CMakeLists.txt:20-56 plus the four jmtx
# translation units of jmtx/CMakeLists.txt:7-18 that provide the CRS container.
set -euo pipefail
REF=${MFV2D_REFERENCE:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
if [ ! -d "$REF/src" ]; then
    echo "build_ref: $REF not present; keeping any prebuilt $OUT" >&2
    exit 0
fi
mkdir -p "$OUT"
PY=${PYTHON:-python3}
PYINC=$($PY -c "import sysconfig; print(sysconfig.get_paths()['include'])")
NPINC=$($PY -c "import numpy; print(numpy.get_include())")
CC=${REF_CC:-/usr/bin/gcc}
TARGET=$OUT/_mfv2d_ref.so
if [ -f "$TARGET" ] && [ -z "${FORCE:-}" ]; then
    echo "build_ref: $TARGET already built"; exit 0
fi
SRCS=$(cd "$REF" && ls src/module.c src/algebra/*.c src/common/*.c src/evaluation/*.c \
        src/fem_space/*.c src/geometry/*.c src/polynomials/*.c \
        jmtx/source/matrix_base.c jmtx/source/modules/double_module.c)
OBJDIR=$(mktemp -d /tmp/mfv2d_ref_obj.XXXXXX)
pids=()
for s in $SRCS; do
    o=$OBJDIR/$(echo "$s" | tr '/' '_').o
    std=-std=gnu17; case "$s" in jmtx/*) std=-std=gnu99;; esac
    $CC $std -O3 -fPIC -fopenmp-simd -DMFV2D_ASSERTS -DNDEBUG -w \
        -I"$PYINC" -I"$NPINC" -I"$REF/jmtx/include" -I"$REF/jmtx/source" \
        -c "$REF/$s" -o "$o" &
    pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
$CC -shared -o "$TARGET" "$OBJDIR"/*.o -lm
rm -rf "$OBJDIR"
echo "build_ref: built $TARGET"
