#!/usr/bin/env bash
# Lists, per object file, the Blackwell-specific SASS mnemonics that prove the kernels are native
# sm_100a code: UTC*MMA (tcgen05.mma), LDTM (tcgen05.ld), UTMALDG/UTMASTG (TMA), LDGMC
# (multimem.ld_reduce), plus register / shared-memory usage from ptxas.
#   tools/sass_summary.sh > profiles/sass_summary.txt
set -euo pipefail
cd "$(dirname "$0")/.."
make -C csrc -j8 >/dev/null
for obj in build/*.o; do
  echo "== $obj"
  cuobjdump -sass "$obj" 2>/dev/null \
    | grep -oE "UTC[A-Z]*MMA[A-Z.0-9_]*|UTCBAR[A-Z.0-9_]*|LDTM[.0-9A-Zx]*|UTMALDG[.0-9A-Z]*|UTMASTG[.0-9A-Z]*|UTMAPF[.0-9A-Z]*|LDGMC[.A-Z0-9x]*|STGMC[.A-Z0-9x]*|SYNCS[.A-Z0-9]*|UTCATOMSWS[.A-Z_]*" \
    | sort | uniq -c | sort -rn | head -20 || true
done
echo
echo "== ptxas resource usage (nvcc -Xptxas -v)"
for src in conv_fprop conv_wgrad conv_wgrad_fused allreduce; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Iinclude -Xptxas -v \
       -c csrc/$src.cu -o /dev/null 2>&1 | grep -E "Compiling entry|Used [0-9]+ registers" \
    | sed -E 's/ptxas info    : //' | paste - - | sed -E 's/Compiling entry function .(_Z[0-9]+|)//' | cut -c1-220
done
