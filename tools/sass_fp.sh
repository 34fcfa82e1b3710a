#!/bin/sh
# SASS instruction mix of one Montgomery product / square per field (evidence for DESIGN.md section 4):
#   sh tools/sass_fp.sh > profiles/sass_fp_mul.txt
set -e
cd "$(dirname "$0")"
T=$(mktemp -d)
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -cubin -o "$T/sass_fp.cubin" sass_fp.cu
echo "# nvcc $(nvcc --version | tail -1); cuobjdump -sass of tools/sass_fp.cu (sm_100a), opcode histogram per kernel"
for k in sass_fq_mul sass_fq_sqr sass_fr_mul sass_fr_sqr; do
    echo "## $k"
    cuobjdump -sass -fun $k "$T/sass_fp.cubin" | grep -E '^\s+/\*[0-9a-f]{4}\*/' | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//' \
        | sed -E 's/^@!?U?P[0-9T]+\s+//' | awk '{print $1}' | sed -E 's/;$//' | sort | uniq -c | sort -rn
done
echo "## excerpt: first 40 instructions of sass_fq_sqr"
cuobjdump -sass -fun sass_fq_sqr "$T/sass_fp.cubin" | grep -E '^\s+/\*[0-9a-f]{4}\*/' | head -60
rm -rf "$T"
