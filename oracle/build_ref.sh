#!/usr/bin/env bash
# oracle/build_ref.sh — TEST INFRASTRUCTURE. Builds the REAL reference (unmodified pebblelib sources
# where they lie under /root/reference + the FFTW 3.3.4 tarball the reference vendors) into
# oracle/_ref/libpebbleref.so. Outputs only under oracle/_ref/ (git-ignored, travels with gpurun).
# Nothing from /root/reference is copied into the tracked tree.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${PEBBLE_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/pebblelib" ]; then
    echo "build_ref: $REF not present (GPU box) — using prebuilt $OUT if any"; exit 0
fi
mkdir -p "$OUT/build"
# 1. FFTW 3.3.4 (the version pebblelib/fftw.h:8 includes), static, PIC, runtime-dispatched SSE2/AVX.
if [ ! -f "$OUT/fftw/lib/libfftw3.a" ]; then
    tar xzf "$REF/External Library Source/fftw-3.3.4.tar.gz" -C "$OUT/build"
    ( cd "$OUT/build/fftw-3.3.4" &&
      ./configure --prefix="$OUT/fftw" --disable-fortran --enable-sse2 --enable-avx --with-pic \
                  --disable-shared --enable-static > ../configure.log 2>&1 &&
      make -j"$(nproc)" > ../make.log 2>&1 && make install > ../install.log 2>&1 )
fi
# 2. include layout so that `#include "../fftw-3.3.4/api/fftw3.h"` (fftw.h:8) resolves through -I.
mkdir -p "$OUT/inc/pebblelib_inc" "$OUT/inc/fftw-3.3.4/api"
cp "$OUT/fftw/include/fftw3.h" "$OUT/inc/fftw-3.3.4/api/fftw3.h"
rm -rf "$OUT/build/fftw-3.3.4"
# 3. the reference sources, compiled in place. -O2 without -march=native: the .so must run on the
#    GPU box's host CPU too. -ffp-contract=off keeps x86-64 SSE2 double semantics (no FMA fusion).
SRCS="cpx db windowfunction fft fftw fftcute fastfir downconvert nco mixer fir decimator"
OBJS=""
CXXFLAGS="-O2 -fPIC -std=c++14 -w -ffp-contract=off -DUSE_FFTW -DPEBBLELIB_LIBRARY \
  -I$HERE/qt_shim -I$OUT/inc/pebblelib_inc -I$REF/pebblelib"
for s in $SRCS; do
    src="$REF/pebblelib/$s.cpp"
    if LC_ALL=C grep -q $'[\x80-\xff]' "$src"; then
        # g++ 13 rejects the typographic quotes/dashes inside `#if 0` prose (db.cpp:70-96): compile a
        # scratch copy with non-ASCII bytes replaced by '?'. Comment text only; no code is touched.
        LC_ALL=C sed $'s/[\x80-\xff]/?/g' "$src" > "$OUT/build/$s.ascii.cpp"
        src="$OUT/build/$s.ascii.cpp"
    fi
    EXTRA=""
    # decimator.h names QVector/QMutex without including them (the Qt build gets them from its precompiled header)
    if [ "$s" = "decimator" ]; then EXTRA="-include QVector -include QMutex -include QDebug"; fi
    g++ $CXXFLAGS $EXTRA -c "$src" -o "$OUT/build/$s.o"
    OBJS="$OBJS $OUT/build/$s.o"
done
g++ $CXXFLAGS -include QVector -include QMutex -c "$HERE/ref_shim.cpp" -o "$OUT/build/ref_shim.o"
g++ -shared -o "$OUT/libpebbleref.so" $OBJS "$OUT/build/ref_shim.o" "$OUT/fftw/lib/libfftw3.a" -lm
echo "build_ref: wrote $OUT/libpebbleref.so"
# 4. BASELINE config 5 (long filter): the same CFastFIR with CONV_FFT_SIZE/CONV_FIR_SIZE 131072/65537. Both are plain
#    #defines (fastfir.cpp:65-66) and FFT clamps sizes at m_maxFFTSize = 65535 (fft.h:21), so this variant is compiled
#    from SCRATCH copies (under _ref/build, deleted afterwards) with exactly those three constants changed by sed; no
#    arithmetic is touched (edits listed in BASELINE.md). Separate .so: the class names are the same.
LD="$OUT/build/long"
rm -rf "$LD"; mkdir -p "$LD"
cp "$REF"/pebblelib/*.h "$LD/"
for s in cpx db windowfunction fft fftw fastfir; do LC_ALL=C sed $'s/[\x80-\xff]/?/g' "$REF/pebblelib/$s.cpp" > "$LD/$s.cpp"; done
sed -i -e 's/^#define CONV_FFT_SIZE 2048/#define CONV_FFT_SIZE 131072/' -e 's/^#define CONV_FIR_SIZE 1025/#define CONV_FIR_SIZE 65537/' "$LD/fastfir.cpp"
sed -i -e 's/m_maxFFTSize = 65535/m_maxFFTSize = 1048576/' "$LD/fft.h"
grep -q 'CONV_FFT_SIZE 131072' "$LD/fastfir.cpp" && grep -q 'CONV_FIR_SIZE 65537' "$LD/fastfir.cpp" && grep -q 'm_maxFFTSize = 1048576' "$LD/fft.h"
LFLAGS="-O2 -fPIC -std=c++14 -w -ffp-contract=off -DUSE_FFTW -DPEBBLELIB_LIBRARY -DREF_SHIM_FASTFIR_ONLY -I$HERE/qt_shim -I$OUT/inc/pebblelib_inc -I$LD"
LOBJS=""
for s in cpx db windowfunction fft fftw fastfir; do g++ $LFLAGS -c "$LD/$s.cpp" -o "$LD/$s.o"; LOBJS="$LOBJS $LD/$s.o"; done
g++ $LFLAGS -include QVector -include QMutex -c "$HERE/ref_shim.cpp" -o "$LD/ref_shim.o"
g++ -shared -o "$OUT/libpebbleref_long.so" $LOBJS "$LD/ref_shim.o" "$OUT/fftw/lib/libfftw3.a" -lm
rm -rf "$LD"
echo "build_ref: wrote $OUT/libpebbleref_long.so"
# 5. Application-side steps: AGC, Demod_AM/NFM/SAM/WFM, SignalSpectrum, SignalStrength, DCRemoval,
#    IQBalance, NoiseBlanker (application/*.cpp) + pebblelib's delayline/iir/fractresampler, UNMODIFIED and compiled where
#    they lie, against the empty Qt stand-ins in qt_shim_app/ (class skeletons only). demod.cpp is not compiled (it includes
#    receiver.h and with it the GUI); ref_app_shim.cpp supplies its two trivial constructors. normalizeIQ: the function
#    bodies deviceinterfacebase.cpp:532-908 are cut into a scratch file (deleted afterwards) that includes devbase_skel.h.
AD="$OUT/build/app"
rm -rf "$AD"; mkdir -p "$AD"
AFLAGS="-O2 -fPIC -std=c++14 -w -ffp-contract=off -DUSE_FFTW -DPEBBLELIB_LIBRARY -include cstring -include QMutex -include QElapsedTimer \
  -I$HERE/qt_shim_app -I$HERE/qt_shim -I$OUT/inc/pebblelib_inc -I$REF/application -I$REF/pebblelib"
AOBJS=""
for s in processstep agc demod/demod_am demod/demod_nfm demod/demod_sam demod/demod_wfm demod/rdsdecode signalstrength \
         signalspectrum dcremoval iqbalance noiseblanker; do
    o="$AD/$(basename $s).o"
    g++ $AFLAGS -c "$REF/application/$s.cpp" -o "$o"; AOBJS="$AOBJS $o"
done
for s in delayline iir fractresampler; do
    g++ $AFLAGS -c "$REF/pebblelib/$s.cpp" -o "$AD/$s.o"; AOBJS="$AOBJS $AD/$s.o"
done
{ echo '#include "devbase_skel.h"'
  awk '/^void DeviceInterfaceBase::normalizeIQ/{on=1} /^void DeviceInterfaceBase::audioProducer/{on=0} on' "$REF/pebblelib/deviceinterfacebase.cpp"
} > "$AD/normalizeiq_cut.cpp"
grep -c 'DeviceInterfaceBase::normalizeIQ' "$AD/normalizeiq_cut.cpp" | grep -qx 11
g++ $AFLAGS -c "$AD/normalizeiq_cut.cpp" -o "$AD/normalizeiq_cut.o"; AOBJS="$AOBJS $AD/normalizeiq_cut.o"
g++ $AFLAGS -c "$HERE/ref_app_shim.cpp" -o "$AD/ref_app_shim.o"
POBJS=""
for s in cpx db windowfunction fft fftw fftcute fir downconvert; do POBJS="$POBJS $OUT/build/$s.o"; done
g++ -shared -o "$OUT/libpebbleref_app.so" $AOBJS "$AD/ref_app_shim.o" $POBJS "$OUT/fftw/lib/libfftw3.a" -lm
rm -rf "$AD"
echo "build_ref: wrote $OUT/libpebbleref_app.so"
# 6. BASELINE config 4 (61.44 MSPS / 15 kHz = 10 stages): MAX_DECSTAGES 10 leaves no NULL sentinel for a 10-stage chain
#    (downconvert.h:23, loop at downconvert.cpp:323). Scratch copy of downconvert.{h,cpp} with that ONE constant
#    raised to 12; no arithmetic touched (edit listed in BASELINE.md). Exports the CDownConvert wrappers only.
DD="$OUT/build/dec12"
rm -rf "$DD"; mkdir -p "$DD"
cp "$REF"/pebblelib/*.h "$DD/"
for s in cpx downconvert; do LC_ALL=C sed $'s/[\x80-\xff]/?/g' "$REF/pebblelib/$s.cpp" > "$DD/$s.cpp"; done
sed -i -e 's/^#define MAX_DECSTAGES 10/#define MAX_DECSTAGES 12/' "$DD/downconvert.h"
grep -q '^#define MAX_DECSTAGES 12' "$DD/downconvert.h"
DFLAGS="-O2 -fPIC -std=c++14 -w -ffp-contract=off -DPEBBLELIB_LIBRARY -DREF_SHIM_DOWNCONVERT_ONLY -I$HERE/qt_shim -I$DD"
g++ $DFLAGS -c "$DD/cpx.cpp" -o "$DD/cpx.o"
g++ $DFLAGS -c "$DD/downconvert.cpp" -o "$DD/downconvert.o"
g++ $DFLAGS -include QVector -include QMutex -c "$HERE/ref_shim.cpp" -o "$DD/ref_shim.o"
g++ -shared -o "$OUT/libpebbleref_dec12.so" "$DD/cpx.o" "$DD/downconvert.o" "$DD/ref_shim.o" -lm
rm -rf "$DD"
echo "build_ref: wrote $OUT/libpebbleref_dec12.so"
# 7. SURVEY a7: the Ooura backend (class CFft, fftcute.cpp) selected the way the reference selects it, -DUSE_FFTCUTE
#    (fft.cpp:45-65). Same unmodified sources, FFT wrappers only.
CD="$OUT/build/cute"
rm -rf "$CD"; mkdir -p "$CD"
CFLAGS_CUTE="-O2 -fPIC -std=c++14 -w -ffp-contract=off -DUSE_FFTCUTE -DPEBBLELIB_LIBRARY -DREF_SHIM_FFT_ONLY -I$HERE/qt_shim -I$OUT/inc/pebblelib_inc -I$REF/pebblelib"
COBJS=""
for s in cpx db windowfunction fft fftcute; do
    src="$REF/pebblelib/$s.cpp"
    if LC_ALL=C grep -q $'[\x80-\xff]' "$src"; then LC_ALL=C sed $'s/[\x80-\xff]/?/g' "$src" > "$CD/$s.ascii.cpp"; src="$CD/$s.ascii.cpp"; fi
    g++ $CFLAGS_CUTE -c "$src" -o "$CD/$s.o"; COBJS="$COBJS $CD/$s.o"
done
g++ $CFLAGS_CUTE -include QVector -include QMutex -c "$HERE/ref_shim.cpp" -o "$CD/ref_shim.o"
g++ -shared -o "$OUT/libpebbleref_cute.so" $COBJS "$CD/ref_shim.o" -lm
rm -rf "$CD"
rm -f "$OUT"/build/*.ascii.cpp
echo "build_ref: wrote $OUT/libpebbleref_cute.so"
