#!/bin/sh
# oracle/build_ref.sh — TEST INFRASTRUCTURE.  Compiles the UNMODIFIED reference sources of the hot path,
# where they lie under /root/reference, against the OpenCV stand-in in oracle/cvshim/ and writes the
# libraries into oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot).
# Never runs the reference's own build system (cmake + OpenCV + Eigen + Pangolin are not in this image).
#
#   libref.so         ORBextractor.cc, system allocator (literal pointer tie-break, src/ORBextractor.cc:927)
#   libref_seq.so     same, operator new = creation-ordered arena (the canonical tie-break)
#   libref_native.so  libref.so built the way the reference's CMakeLists.txt:10-13 builds it
#                     (-O3 -march=native there; here -march=x86-64-v3 = AVX2+FMA, because the library is
#                     built in this container and must run on the GPU box's unknown host CPU: host FMA
#                     contraction allowed) — the timed CPU baseline
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${ORB_REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -f "$REF/src/ORBextractor.cc" ]; then
    echo "build_ref.sh: $REF/src/ORBextractor.cc not found (expected on the GPU box: prebuilt oracle/_ref is used)"
    exit 0
fi
mkdir -p "$OUT"
# the system compiler, not $CXX: this image exports CXX=/opt/gcc/bin/g++, a wrapper that links libstdc++ STATICALLY; a private
# copy of the iostream / locale state inside a dlopen'ed library crashes once the process also holds the shared libstdc++
# (numpy and torch load it).  ORB_REF_CXX / ORB_REF_CC override.
CXX="${ORB_REF_CXX:-g++}"
CC="${ORB_REF_CC:-gcc}"
[ -x /usr/bin/g++ ] && [ -z "$ORB_REF_CXX" ] && CXX=/usr/bin/g++
[ -x /usr/bin/gcc ] && [ -z "$ORB_REF_CC" ] && CC=/usr/bin/gcc
COMMON="-O3 -fPIC -w -I$HERE/cvshim -I$REF/include -I$HERE"
CANON="-ffp-contract=off"
LINK="-shared -Wl,-Bsymbolic -lm -lpthread"

$CC -O3 -fPIC -std=c11 $CANON -c "$HERE/orc_prims.c" -o "$OUT/orc_prims.o"
$CC -O3 -fPIC -std=c11 -march=x86-64-v3 -c "$HERE/orc_prims.c" -o "$OUT/orc_prims_native.o"

$CXX -std=c++14 $COMMON $CANON -c "$REF/src/ORBextractor.cc" -o "$OUT/ORBextractor.o"
$CXX -std=c++14 $COMMON $CANON -c "$HERE/ref_wrap.cc" -o "$OUT/ref_wrap.o"
$CXX -o "$OUT/libref.so" "$OUT/ORBextractor.o" "$OUT/ref_wrap.o" "$OUT/orc_prims.o" $LINK

$CXX -std=c++14 $COMMON $CANON -DREF_SEQ_ALLOC -c "$HERE/ref_wrap.cc" -o "$OUT/ref_wrap_seq.o"
$CXX -o "$OUT/libref_seq.so" "$OUT/ORBextractor.o" "$OUT/ref_wrap_seq.o" "$OUT/orc_prims.o" $LINK

$CXX -std=c++14 $COMMON -march=x86-64-v3 -c "$REF/src/ORBextractor.cc" -o "$OUT/ORBextractor_native.o"
$CXX -std=c++14 $COMMON -march=x86-64-v3 -c "$HERE/ref_wrap.cc" -o "$OUT/ref_wrap_native.o"
$CXX -o "$OUT/libref_native.so" "$OUT/ORBextractor_native.o" "$OUT/ref_wrap_native.o" "$OUT/orc_prims_native.o" $LINK

# libref_match.so: the UNMODIFIED src/ORBmatcher.cc over oracle/refshim's Frame / KeyFrame / MapPoint / camera doubles
# (force-included in front of it; they define the include guards of the real headers) and the restated grid of orc_match.c
$CC -O3 -fPIC -std=c11 $CANON -c "$HERE/orc_match.c" -o "$OUT/orc_match.o"
MATCH="-I$HERE/refshim -I$REF -include $HERE/refshim/ref_stubs.h"
$CXX -std=c++14 $COMMON $CANON $MATCH -c "$REF/src/ORBmatcher.cc" -o "$OUT/ORBmatcher.o"
$CXX -std=c++14 $COMMON $CANON $MATCH -c "$HERE/ref_match_wrap.cc" -o "$OUT/ref_match_wrap.o"
$CXX -o "$OUT/libref_match.so" "$OUT/ORBmatcher.o" "$OUT/ref_match_wrap.o" "$OUT/orc_match.o" "$OUT/orc_prims.o" $LINK

# libref_bow.so: the reference's vendored, UNMODIFIED DBoW2 (Thirdparty/DBoW2) - ORBVocabulary::loadFromTextFile / transform
DBOW="$REF/Thirdparty/DBoW2"
if [ -f "$DBOW/DBoW2/TemplatedVocabulary.h" ]; then
    BOW="-O3 -fPIC -w $CANON -I$HERE/cvshim -I$HERE/boostshim -I$DBOW"
    BOWOBJ=""
    for f in DBoW2/BowVector DBoW2/FORB DBoW2/FeatureVector DBoW2/ScoringObject DUtils/Random DUtils/Timestamp; do
        o="$OUT/dbow_$(basename $f).o"
        $CXX -std=c++14 $BOW -c "$DBOW/$f.cpp" -o "$o"
        BOWOBJ="$BOWOBJ $o"
    done
    $CXX -std=c++14 $BOW -c "$HERE/ref_bow_wrap.cc" -o "$OUT/ref_bow_wrap.o"
    $CXX -o "$OUT/libref_bow.so" $BOWOBJ "$OUT/ref_bow_wrap.o" "$OUT/orc_prims.o" $LINK

    # libref_frame.so: the UNMODIFIED src/Frame.cc with the real include/Frame.h, ORBextractor.cc and the vendored DBoW2; doubles of
    # the IMU / camera / map-point headers in oracle/refshim_frame (force-included); the oracle's C primitives behind the OpenCV
    # stand-in's undistortPoints / BFMatcher / FAST / resize / GaussianBlur
    FRAME="-I$HERE/boostshim -I$HERE/refshim_frame -I$REF -I$REF/include/CameraModels -I$DBOW -include $HERE/refshim_frame/ref_frame_stubs.h"
    $CXX -std=c++14 $COMMON $CANON $FRAME -c "$REF/src/Frame.cc" -o "$OUT/Frame.o"
    $CXX -std=c++14 $COMMON $CANON $FRAME -c "$HERE/ref_frame_wrap.cc" -o "$OUT/ref_frame_wrap.o"
    ORCOBJ=""
    for f in orc_extractor orc_octree orc_detect orc_prepare; do
        $CC -O3 -fPIC -std=c11 $CANON -c "$HERE/$f.c" -o "$OUT/$f.o"
        ORCOBJ="$ORCOBJ $OUT/$f.o"
    done
    $CXX -o "$OUT/libref_frame.so" "$OUT/Frame.o" "$OUT/ref_frame_wrap.o" "$OUT/ORBextractor.o" $BOWOBJ \
        "$OUT/orc_match.o" "$OUT/orc_prims.o" $ORCOBJ $LINK
fi

# libref_mp.so: the UNMODIFIED src/MapPoint.cc with the real include/MapPoint.h over doubles of KeyFrame / Frame / Map
MP="-I$HERE/boostshim -I$HERE/refshim_mp -I$REF -include $HERE/refshim_mp/ref_mp_stubs.h"
$CXX -std=c++14 $COMMON $CANON $MP -c "$REF/src/MapPoint.cc" -o "$OUT/MapPoint.o"
$CXX -std=c++14 $COMMON $CANON $MP -c "$HERE/ref_mp_wrap.cc" -o "$OUT/ref_mp_wrap.o"
$CXX -o "$OUT/libref_mp.so" "$OUT/MapPoint.o" "$OUT/ref_mp_wrap.o" "$OUT/orc_match.o" "$OUT/orc_prims.o" $LINK

rm -f "$OUT"/*.o
sha256sum "$REF/src/ORBextractor.cc" "$REF/include/ORBextractor.h" "$REF/src/ORBmatcher.cc" "$REF/include/ORBmatcher.h" "$REF/src/Frame.cc" "$REF/include/Frame.h" "$REF/src/MapPoint.cc" "$REF/include/MapPoint.h" "$REF/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h" "$REF/Thirdparty/DBoW2/DBoW2/FORB.cpp" > "$OUT/SOURCES.sha256" 2>/dev/null || true
echo "build_ref.sh: built $(ls "$OUT"/*.so | wc -l) libraries in $OUT"
