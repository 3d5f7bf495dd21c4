# Build an experiment variant of the library next to the shipped one:  tools/build_variant.sh <name> [-DFLAG=...]...
#   -> _build/ab/<name>.so   (select at run time with GASMA_B200_LIB; A/B scripts: tools/ab_*.sh)
NAME=$1; shift
mkdir -p _build/ab
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared \
  -Xlinker -Bsymbolic -Xlinker --exclude-libs,ALL -ldl -lpthread --fmad=true "$@" \
  -I include -o _build/ab/$NAME.so approximate-string-matching_b200/csrc/gasma_b200.cu
