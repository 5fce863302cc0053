#!/bin/bash
# INTEGRATION.md "Level 0": the reference's sdrx.cpp compiles, unmodified, against the drop-in class headers.
# sdrx.cpp includes its DSP classes with quotes ("msd.hpp"), which resolve next to the including file, so the swap is
# a swap of FILES: a shadow source directory holds symlinks to every reference source except iqsample / msd / ds /
# fir / agc.hpp, which point at include/sdrx_b200/.  popt, fftw3f, alsa and sigc++ are absent from this image:
# declarations-only stand-ins under tests/cpp/stubs/ make `g++ -fsyntax-only` possible (nothing is linked or run).
# usage: check_sdrx_syntax.sh [reference_src_dir]      exit 0 = compiles
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(cd "$HERE/../.." && pwd)
REF=${1:-/root/reference/src}
SH=$HERE/_build/shadow
rm -rf "$SH"; mkdir -p "$SH"
for f in "$REF"/*; do ln -s "$f" "$SH/$(basename "$f")"; done
for h in iqsample.hpp msd.hpp ds.hpp fir.hpp agc.hpp; do rm -f "$SH/$h"; ln -s "$ROOT/include/sdrx_b200/$h" "$SH/$h"; done
for variant in "" "-DUSE_THREAD_POOL"; do
    g++ -std=c++20 -fsyntax-only -Wall $variant -I"$SH" -I"$ROOT/include" -I"$HERE/stubs" "$SH/sdrx.cpp"
    echo "sdrx.cpp $variant: syntax OK against include/sdrx_b200/{iqsample,msd,ds,fir,agc}.hpp"
done
