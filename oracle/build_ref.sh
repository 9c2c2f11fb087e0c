#!/bin/bash
# Build the REAL reference (Thomi6York/pbrt-v2-skin) hot path as a CPU checker.
# TEST INFRASTRUCTURE ONLY -- see oracle/README.md.  Nothing in the product path uses it.
#
#   input : /root/reference/src  (read-only; only present in the build container)
#   output: oracle/_ref/ref_driver   (git-ignored binary; travels to the GPU box with gpurun)
#
# The reference is MSVC-only (README.md:42-43): its sources are copied to a mktemp
# directory, patched there by refbuild/patch_ref.py (spelling fixes only),
# compiled file by file with g++ (the reference's own Makefile is NOT run: it needs flex,
# bison and OpenEXR), linked with refbuild/ref_driver.cpp, and the temp copy is deleted.
# No reference source is ever written into this repository.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${PBRT_SKIN_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src/core" ]; then
  echo "build_ref: $REF/src not present (GPU box?) -- keeping prebuilt $OUT" >&2
  exit 0
fi
STAMP="$OUT/.stamp"
SIG="$(cat "$HERE/refbuild/ref_driver.cpp" "$HERE/refbuild/patch_ref.py" "$0" | sha1sum | cut -d' ' -f1)"
if [ -x "$OUT/ref_driver" ] && [ -f "$STAMP" ] && [ "$(cat "$STAMP")" = "$SIG" ]; then
  echo "build_ref: up to date"; exit 0
fi
mkdir -p "$OUT"
TMP="$(mktemp -d /tmp/pbrtskin_ref.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
cp -r "$REF/src" "$TMP/src"
chmod -R u+w "$TMP/src"
python3 "$HERE/refbuild/patch_ref.py" "$TMP/src"
cd "$TMP/src"
mkdir -p objs
CXXF="-std=gnu++11 -fpermissive -include limits -D__declspec(x)= -O2 -m64 -DNDEBUG -DPBRT_PROBES_NONE -w"
compile_one() {
  f="$1"; o="objs/$(echo "${f%.*}" | tr '/' '_').o"
  case "$f" in
    multipole/MultipoleProfileCalculator/*.cpp) g++ $CXXF -Imultipole/libkissfft -c "$f" -o "$o" ;;
    *.c) gcc -O2 -m64 -I. -Icore -Imultipole/libkissfft -DNDEBUG -w -c "$f" -o "$o" ;;
    *) g++ $CXXF -I. -Icore -c "$f" -o "$o" ;;
  esac
}
export -f compile_one; export CXXF
ls core/*.cpp core/targa.c accelerators/*.cpp cameras/*.cpp film/*.cpp filters/*.cpp integrators/*.cpp \
   lights/*.cpp materials/*.cpp renderers/*.cpp samplers/*.cpp shapes/*.cpp textures/*.cpp volumes/*.cpp \
   multipole/MultipoleProfileCalculator/*.cpp multipole/libkissfft/kiss_fft.c \
   multipole/libkissfft/tools/kiss_fftr.c multipole/libkissfft/tools/kiss_fftnd.c \
   multipole/libkissfft/tools/kiss_fftndr.c \
  | grep -v -e 'core/parser.cpp' -e 'core/api.cpp' -e 'dllmain' -e 'stdafx.cpp' \
  | xargs -P "$(nproc)" -I{} bash -c 'compile_one {}'
g++ $CXXF -I. -Icore -c "$HERE/refbuild/ref_driver.cpp" -o objs/ref_driver.o
g++ -O2 -o "$OUT/ref_driver" objs/*.o -lpthread
echo "$SIG" > "$STAMP"
echo "build_ref: built $OUT/ref_driver"
