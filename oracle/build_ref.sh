#!/bin/bash
# Builds oracle/_ref/libsgz_oracle_ref.so: the REFERENCE's own Spectrum transform code behind the oracle's C entry points.
#
#   * copies Source/Spectrum/{TransformDSP.inl,TransformPair.h,TransformConstant.h} from the reference checkout
#     (default /root/reference; read-only) into a scratch directory under /tmp that is removed afterwards - no
#     reference source is ever stored here; only the built library lands in oracle/_ref/ (git-ignored);
#   * puts the stand-in headers of oracle/ref_stubs/ (cpl/*, Signalizer.h, SpectrumParameters.h) around them: the
#     reference's numeric library cpl is absent from the checkout, so its primitives forward to the oracle's "cpl
#     decisions" (FFT, interpolation filters, window, resonator);
#   * compiles oracle/sgz_oracle_capi.cpp with -DSGZ_ORACLE_REF (ref_adapter.hpp), same flags as the parity build of
#     the restated oracle (-O2 -ffp-contract=off).
# Without the reference checkout (the GPU box) it does nothing and keeps whatever oracle/_ref/ already holds.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
ref="${SGZ_REFERENCE:-/root/reference}/Source/Spectrum"
out="$here/_ref"
if [ ! -f "$ref/TransformDSP.inl" ]; then
	echo "build_ref: no reference checkout at $ref, keeping $out as it is"
	exit 0
fi
mkdir -p "$out"
stamp="$out/libsgz_oracle_ref.so"
if [ -f "$stamp" ] && [ -z "$(find "$here"/*.hpp "$here"/*.cpp "$here/ref_stubs" "$ref/TransformDSP.inl" "$ref/TransformPair.h" "$ref/TransformConstant.h" -newer "$stamp" 2>/dev/null)" ]; then
	exit 0
fi
# scratch tree: stand-in headers + the three reference files side by side, as their relative includes expect
src="$(mktemp -d)"
trap 'rm -rf "$src"' EXIT
mkdir -p "$src/Spectrum"
cp -r "$here/ref_stubs/cpl" "$src/"
cp "$here/ref_stubs/Signalizer.h" "$src/"
cp "$here/ref_stubs/Spectrum/SpectrumParameters.h" "$src/Spectrum/"
for f in TransformDSP.inl TransformPair.h TransformConstant.h; do cp "$ref/$f" "$src/Spectrum/$f"; done
flags="-std=c++17 -shared -fPIC -pthread -ffp-contract=off -Wall -Wno-unused-variable -Wno-unused-but-set-variable -Wno-sign-compare -Wno-unused-local-typedefs -Wno-switch -Wno-class-memaccess -DSGZ_ORACLE_REF"
# parity build (same flags as libsgz_oracle.so) and the timed build (same flags as libsgz_oracle_fast.so)
${CXX:-g++} $flags -O2 -I"$src" -I"$here" -o "$stamp" "$here/sgz_oracle_capi.cpp"
${CXX:-g++} $flags -O3 -mavx2 -mfma -I"$src" -I"$here" -o "$out/libsgz_oracle_ref_fast.so" "$here/sgz_oracle_capi.cpp"
echo "build_ref: built $stamp and its -O3 twin"
