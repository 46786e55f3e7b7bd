#!/bin/sh
# TEST INFRASTRUCTURE ONLY. Builds the UNMODIFIED reference hot path
# (/root/reference/impl/dataCompression.c) into oracle/_ref/libdcref_<eb>.so,
# one shared object per absErrorBound (the bound is a compile-time macro,
# impl/dataCompression.h:5; rewritten with the reference's own sed recipe,
# impl/set-parameter.sh:8). No reference source is copied into the repo: the
# .c is fed to gcc on stdin from where it lies; the sed-rewritten header is a
# temporary build product that is deleted again after the compile.
# Usage: oracle/build_ref.sh [eb ...]   (default: 0.1 0.01 0.001 0.0001 0.00001 0.000001)
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
REF=${DC_REFERENCE_DIR:-/root/reference}/impl
OUT=$HERE/_ref
[ -f "$REF/dataCompression.c" ] || { echo "build_ref: no reference at $REF (skipped)"; exit 0; }
mkdir -p "$OUT"
EBS="$@"
[ -n "$EBS" ] || EBS="0.1 0.01 0.001 0.0001 0.00001 0.000001"
for eb in $EBS; do
  so="$OUT/libdcref_$eb.so"
  [ -f "$so" ] && [ "$so" -nt "$REF/dataCompression.c" ] && continue
  tmp="$OUT/.hdr_$eb.$$"
  mkdir -p "$tmp"
  sed "s/#define absErrorBound[[:space:]][[:space:]]*[^ \t][^ \t]*/#define absErrorBound\t$eb/" \
      "$REF/dataCompression.h" > "$tmp/dataCompression.h"
  # quote-includes resolve against the cwd when the source comes from stdin
  (cd "$tmp" && gcc -x c -O3 -ffp-contract=off -fPIC -shared -w \
      -I"$HERE/stub" -I"$REF" -o "$so" - -lm -lz < "$REF/dataCompression.c")
  rm -rf "$tmp"
  echo "built $so"
done
