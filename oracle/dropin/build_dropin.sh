#!/bin/sh
# TEST INFRASTRUCTURE ONLY.  Links the reference's own MPI drivers against libdcb200.so exactly as INTEGRATION.md section 2
# prescribes (objcopy --localize-symbol on the reference object, -ldcb200 before it), with the single-process MPI stand-in of
# this directory instead of an MPI installation.  Sources are compiled where they lie under /root/reference; outputs go to
# oracle/_ref/dropin/ (git-ignored, travels to the GPU box):
#   pingpong_dcb, bmt_dcb   the drivers with the 14 hot symbols bound to libdcb200.so
#   pingpong_ref            the same driver linked against the unmodified reference object (validates the stand-in)
#   libhimeno_ref.so        impl/himenoBMTxps.c for ONE rank (param.h by the reference's own paramset.sh) + accessors
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
REPO=$(cd "$HERE/../.." && pwd)
REF=${DC_REFERENCE_DIR:-/root/reference}/impl
OUT=$REPO/oracle/_ref/dropin
DCB=$REPO/data_compression_b200
[ -f "$REF/pingpong.c" ] || { echo "build_dropin: no reference at $REF (skipped)"; exit 0; }
mkdir -p "$OUT"
CC=${CC:-gcc}
HOT="myCompress_bitwise myDecompress_bitwise myCompress_bitwise_np myDecompress_bitwise_np myCompress_bitwise_op myDecompress_bitwise_op myCompress_bitwise_mask myDecompress_bitwise_mask toSmallDataset_float med_dataset_float do_crc32 floattostr strtofloat to_absErrorBound_binary"
echo "$HOT" | tr ' ' '\n' > "$OUT/hot_symbols.txt"
CFLAGS="-O3 -ffp-contract=off -w -I$HERE -I$REF"
$CC $CFLAGS -c "$HERE/mpistub.c" -o "$OUT/mpistub.o"
$CC $CFLAGS -c "$REF/dataCompression.c" -o "$OUT/dataCompression.o"
$CC $CFLAGS -c "$REF/pingpong.c" -o "$OUT/pingpong.o"
$CC $CFLAGS -c "$REF/himenoBMTxps.c" -o "$OUT/himenoBMTxps.o"
# the object keeps using its own copies internally, the drivers bind to the library
LOCALIZE=""
for s in $HOT; do LOCALIZE="$LOCALIZE --localize-symbol=$s"; done
objcopy $LOCALIZE "$OUT/dataCompression.o" "$OUT/dataCompression_cold.o"
$CC -o "$OUT/pingpong_dcb" "$OUT/pingpong.o" -L"$DCB" -ldcb200 "$OUT/dataCompression_cold.o" "$OUT/mpistub.o" -lm -lz -Wl,-rpath,'$ORIGIN/../../../data_compression_b200'
$CC -o "$OUT/bmt_dcb" "$OUT/himenoBMTxps.o" -L"$DCB" -ldcb200 "$OUT/dataCompression_cold.o" "$OUT/mpistub.o" -lm -lz -Wl,-rpath,'$ORIGIN/../../../data_compression_b200'
$CC -o "$OUT/pingpong_ref" "$OUT/pingpong.o" "$OUT/dataCompression.o" "$OUT/mpistub.o" -lm -lz
# one-rank Himeno (size S, 1 x 1 x 1) as a shared object: the reference's source on stdin, so that "param.h" resolves in cwd
tmp="$OUT/.himeno.$$"
mkdir -p "$tmp"
(cd "$tmp" && sh "$REF/paramset.sh" S 1 1 1 >/dev/null && cp "$HERE/mpi.h" . &&
 $CC -x c -O3 -ffp-contract=off -fPIC -w -I. -I"$REF" -c -o dc_S.o - < "$REF/dataCompression.c" &&
 $CC -O3 -fPIC -w -I. -c -o mpistub_pic.o "$HERE/mpistub.c" &&
 (echo '#define main himeno_main'; cat "$REF/himenoBMTxps.c"; echo; echo '#undef main'; cat "$HERE/himeno_access.c") |
 $CC -x c -O3 -ffp-contract=off -fPIC -w -I. -I"$REF" -c -o himeno_S.o - &&
 $CC -shared -o "$OUT/libhimeno_ref.so" himeno_S.o dc_S.o mpistub_pic.o -lm -lz)
rm -rf "$tmp"
echo "built $OUT"
