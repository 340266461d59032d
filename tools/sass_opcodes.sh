#!/bin/bash
# Counts of the Blackwell-specific SASS opcodes per kernel of the built library (evidence that the tensor-core and TMA
# paths are what is compiled in): UTCIMMA (tcgen05.mma), LDTM (tcgen05.ld), UTCBAR (tcgen05.commit), UBLKCP / UTMALDG
# (bulk / tensor TMA copies), SYNCS (mbarrier), plus POPC and IMMA for the matching engines.
#   tools/sass_opcodes.sh > profiles/sass_opcodes.txt
SO=${1:-visual-slam_b200/orbslam2/libvslam_b200.so}
echo "# cuobjdump -sass $SO  (nvcc $(nvcc --version | grep release | sed 's/.*release //'))"
cuobjdump -sass "$SO" | awk '
/Function : / { fn=$3 }
{ for (i = 1; i <= NF; ++i) {
    op=$i; sub(/\..*/, "", op);
    if (op ~ /^(UTCIMMA|UTCHMMA|LDTM|STTM|UTCBAR|UBLKCP|UTMALDG|UTMASTG|SYNCS|POPC|IMMA|VIMNMX3|IDP)$/) c[fn" "op]++ } }
END { for (k in c) print k, c[k] }' | sort | while read fn op n; do printf "%-8s %5d  %s\n" "$op" "$n" "$(echo $fn | c++filt | cut -c1-110)"; done
