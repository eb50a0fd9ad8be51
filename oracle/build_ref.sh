#!/usr/bin/env bash
# TEST INFRASTRUCTURE.  Builds the UNMODIFIED reference (eclipse-sumo/sumo) `sumo` and `netgenerate`
# binaries from the sources where they lie under /root/reference/src, into oracle/_ref/ (git-ignored).
# The reference's own build system is not run.  Its one missing hard dependency, Xerces-C
# (CMakeLists.txt:476), is satisfied by the expat-backed header shim in oracle/xerces_shim/.
# Flags: -O2 -ffp-contract=off and no -march, so the oracle is strict non-FMA IEEE fp64
# (the reference sets no FP flags itself, CMakeLists.txt:147-157).
#
# usage: oracle/build_ref.sh [sumo|netgenerate|duarouter|flatten|sumo_b200|all]   (default: all; sumo_b200 needs sumo_b200/libsumob200.so)
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SUMO_REFERENCE:-/root/reference}"
SRC="$REF/src"
OUT="$HERE/_ref"
WHAT="${1:-all}"
JOBS="${JOBS:-$(nproc)}"

if [ ! -d "$SRC" ]; then
    echo "build_ref: $SRC not present (GPU box?) -- using prebuilt files in $OUT" >&2
    exit 0
fi
mkdir -p "$OUT/gen" "$OUT/obj"

# --- generated headers (what cmake's configure_file / version.py would write) -------------------
cat > "$OUT/gen/config.h" <<'EOF'
/* hand-written equivalent of src/config.h.cmake with every optional dependency off */
#pragma once
#define FOX_CONSTRUCTOR(classname) classname() {}
#define HAVE_ENABLED "oracle-build (expat shim)"
#define HAVE_VERSION_H
#define HAVE_ZLIB   /* zlib.h is in the image; without it utils/iodevices/OutputDevice_File.cpp:68 does not compile */
#define NUMERICAL_EPS 0.001
#define POSITION_EPS 0.1
#define SPEED_EPS 0.01
#define XERCES3_SIZE_t XMLSize_t
EOF
cat > "$OUT/gen/version.h" <<'EOF'
#pragma once
#define VERSION_STRING "1.27.1-oracle"
EOF

# netimport/typemap.h is produced by the reference's own generator script (netimport/CMakeLists.txt:51)
mkdir -p "$OUT/gen/netimport" "$OUT/gen/polyconvert"
python3 "$REF/tools/build_config/typemap.py" "$OUT/gen"

CXXFLAGS="-std=c++17 -O2 -ffp-contract=off -fPIC -w -DNDEBUG -I$OUT/gen -I$HERE/xerces_shim -I$SRC"

srcs_in() { # all .cpp below the given dirs (relative to $SRC), minus the ones the reference's
            # CMake lists only add when an optional dependency was found (JuPedSim, Parquet)
    for d in "$@"; do find "$SRC/$d" -name '*.cpp' | grep -v -E 'MSPModel_JuPedSim|ParquetFormatter'; done
}

COMMON="$(srcs_in utils/common utils/distribution utils/geom utils/handlers utils/importio utils/iodevices \
           utils/options utils/shapes utils/traction_wire utils/xml foreign/tcpip)"
VEHICLE="$(srcs_in utils/emissions utils/vehicle foreign/PHEMlight/cpp foreign/PHEMlight/V5/cpp)"
SIM="$(srcs_in microsim mesosim netload traci-server) $(ls $SRC/libsumo/*.cpp | grep -v '/GUI.cpp')"
NET="$(srcs_in netbuild netwrite netimport foreign/eulerspiral) $(ls $SRC/netgen/*.cpp)"
DUA="$(srcs_in router) $(ls $SRC/duarouter/*.cpp)"

# overlay copy of MSNet.cpp: declaration + one inserted line at the top of MSNet::simulationStep (src/microsim/MSNet.cpp:799)
sed -e 's|^MSNet::simulationStep(const bool onlyMove) {$|MSNet::simulationStep(const bool onlyMove) {\n    if (sumob200_hook_step(this)) {\n        return;\n    }|' \
    -e '0,/^#include <config.h>/s||#include <config.h>\nclass MSNet;\nbool sumob200_hook_step(MSNet* net);   // integration/sumo_b200_hook.cpp|' \
    "$SRC/microsim/MSNet.cpp" > "$OUT/gen/MSNet_b200.cpp"
grep -q 'sumob200_hook_step(this)' "$OUT/gen/MSNet_b200.cpp" || { echo "build_ref: could not patch MSNet::simulationStep" >&2; exit 1; }

MK="$OUT/Makefile"
{
    echo "CXX=g++"
    echo "CXXFLAGS=$CXXFLAGS"
    echo "OBJ=$OUT/obj"
    emit() { # name, main-less source list
        local name="$1"; shift
        local objs=""
        for s in "$@"; do
            local o="\$(OBJ)/$(echo "${s#$SRC/}" | tr '/' '_' | sed 's/\.cpp$/.o/')"
            objs="$objs $o"
            echo "$o: $s"
            printf '\t@echo CXX %s\n\t@$(CXX) $(CXXFLAGS) -c $< -o $@\n' "${s#$SRC/}"
        done
        echo "${name}_OBJS=$objs"
    }
    emit COMMON $COMMON
    emit VEHICLE $VEHICLE
    emit SIM $SIM $SRC/sumo_main.cpp
    emit NET $NET
    emit DUA $DUA
    echo "$OUT/sumo: \$(SIM_OBJS) \$(VEHICLE_OBJS) \$(COMMON_OBJS)"
    printf '\t@echo LINK $@\n\t@$(CXX) -o $@ $^ -l:libexpat.a -l:libz.a -lpthread -ldl\n'
    echo "$OUT/netgenerate: \$(NET_OBJS) \$(VEHICLE_OBJS) \$(COMMON_OBJS)"
    printf '\t@echo LINK $@\n\t@$(CXX) -o $@ $^ -l:libexpat.a -l:libz.a -lpthread -ldl\n'
    echo "$OUT/duarouter: \$(DUA_OBJS) \$(VEHICLE_OBJS) \$(COMMON_OBJS)"
    printf '\t@echo LINK $@\n\t@$(CXX) -o $@ $^ -l:libexpat.a -l:libz.a -lpthread -ldl\n'
    # integration/sumo_flatten.cpp = the reference-side flattening shim + parity trace recorder; it replaces sumo_main.o
    echo "\$(OBJ)/sumo_flatten.o: $HERE/../integration/sumo_flatten.cpp $HERE/../include/sumob200_flatfile.h"
    printf '\t@echo CXX integration/sumo_flatten.cpp\n\t@$(CXX) $(CXXFLAGS) -c $< -o $@\n'
    echo "$OUT/sumo_flatten: \$(OBJ)/sumo_flatten.o \$(filter-out %sumo_main.o,\$(SIM_OBJS)) \$(VEHICLE_OBJS) \$(COMMON_OBJS)"
    printf '\t@echo LINK $@\n\t@$(CXX) -o $@ $^ -l:libexpat.a -l:libz.a -lpthread -ldl\n'
    # sumo_b200 = the reference's sumo driving libsumob200.so (integration/sumo_b200_hook.cpp): MSNet.cpp compiled from an overlay copy with
    # the one-line hook of INTEGRATION.md inserted by sed (generated below, never committed); needs sumo_b200/libsumob200.so
    echo "\$(OBJ)/MSNet_b200.o: $OUT/gen/MSNet_b200.cpp"
    printf '\t@echo CXX gen/MSNet_b200.cpp\n\t@$(CXX) $(CXXFLAGS) -I%s -c $< -o $@\n' "$SRC/microsim"
    echo "\$(OBJ)/sumo_b200_hook.o: $HERE/../integration/sumo_b200_hook.cpp $HERE/../integration/sumo_flatten.cpp $HERE/../include/sumob200_flatfile.h $HERE/../include/sumob200.h"
    printf '\t@echo CXX integration/sumo_b200_hook.cpp\n\t@$(CXX) $(CXXFLAGS) -c $< -o $@\n'
    echo "$OUT/sumo_b200: \$(OBJ)/sumo_b200_hook.o \$(OBJ)/MSNet_b200.o \$(filter-out %sumo_main.o %microsim_MSNet.o,\$(SIM_OBJS)) \$(VEHICLE_OBJS) \$(COMMON_OBJS)"
    printf '\t@echo LINK $@\n\t@$(CXX) -o $@ $^ -L%s -lsumob200 -Wl,-rpath,\\$$ORIGIN/../../sumo_b200 -l:libexpat.a -l:libz.a -lpthread -ldl\n' "$HERE/../sumo_b200"
    echo "sumo_b200: $OUT/sumo_b200"
    echo "flatten: $OUT/sumo_flatten"
    echo "sumo: $OUT/sumo"
    echo "netgenerate: $OUT/netgenerate"
    echo "duarouter: $OUT/duarouter"
    echo "all: sumo netgenerate flatten"
} > "$MK"

make -k -f "$MK" -j"$JOBS" "$WHAT"
