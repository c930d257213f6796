#!/bin/bash
# oracle/build_gpu.sh -- builds the reference tree WITH the product's integration patches applied, linked against
# haploclique_b200/libhcgpu.so: oracle/_ref/libhcref_gpu.so (test face) and oracle/_ref/haploclique_gpu (the CLI).
#   integration/CLEVER_rows.patch, BronKerbosch_rows.patch   the finders' edge loops consume adjacency rows
#   integration/CliqueCollector_batch.patch                  the collector hands a round's cliques to the GPU merge (N2)
# The patched copies of the reference's src/ live in a temporary directory for the duration of the compile only: no
# reference source stays in the repo, outputs go to oracle/_ref/ (git-ignored). The unmodified reference objects stay in
# libhcref.so / haploclique_ref. Called by `make -C oracle gpu`; arguments: REF OUT "<REFFLAGS>" "<BAM_OBJS>" "<docopt.o>"
set -euo pipefail
REF="$1"; OUT="$2"; REFFLAGS="$3"; BAM_OBJS="$4"; DOCOPT_O="$5"
HERE="$(cd "$(dirname "$0")" && pwd)"
cd "$HERE"
T="$(mktemp -d)"
trap 'rm -rf "$T"' EXIT
cp "$REF"/src/*.cpp "$REF"/src/*.h "$T"/
chmod u+w "$T"/*
for p in CLEVER_rows BronKerbosch_rows CliqueCollector_batch; do patch -s -p1 -d "$T" < ../integration/$p.patch; done
cp "$OUT"/version.h "$T"/version.h
OBJ="$OUT/obj/gpu"
mkdir -p "$OBJ"
FLAGS="-I$T $REFFLAGS -I../integration -I../include -DHC_WITH_GPU_ADAPTER"
compile() { g++ $FLAGS -c "$1" -o "$2"; }
export -f compile; export FLAGS
jobs=()
for n in AlignmentRecord ShortDnaSequence NewEdgeCalculator QuasispeciesEdgeCalculator CLEVER BronKerbosch Clique GaussianEdgeCalculator; do
    jobs+=("$T/$n.cpp $OBJ/$n.o")
done
jobs+=("../integration/GpuEdgeCalculator.cpp $OBJ/GpuEdgeCalculator.o" "../integration/GpuSuperReadBuilder.cpp $OBJ/GpuSuperReadBuilder.o" "../integration/GpuBamLoader.cpp $OBJ/GpuBamLoader.o"
       "ref_api.cpp $OBJ/ref_api_gpu.o" "../integration/haploclique_gpu_main.cpp $OBJ/gpu_main.o")
printf '%s\n' "${jobs[@]}" | xargs -P 8 -L 1 bash -c 'compile "$0" "$1"'
REF_OBJS=""
for n in AlignmentRecord ShortDnaSequence NewEdgeCalculator QuasispeciesEdgeCalculator CLEVER BronKerbosch Clique GaussianEdgeCalculator GpuEdgeCalculator GpuSuperReadBuilder GpuBamLoader; do REF_OBJS="$REF_OBJS $OBJ/$n.o"; done
LINK="-L../haploclique_b200 -lhcgpu -Wl,-rpath,\$ORIGIN/../../haploclique_b200 -lz -lpthread"
g++ -shared -o "$OUT/libhcref_gpu.so" "$OBJ/ref_api_gpu.o" $REF_OBJS $DOCOPT_O $BAM_OBJS $LINK
g++ -o "$OUT/haploclique_gpu" "$OBJ/gpu_main.o" $REF_OBJS $DOCOPT_O $BAM_OBJS $LINK
