#!/bin/bash
# static SASS opcode histogram of the single-surface probe kernels (scratch/one.cu)
cd /root/repo/optiland_pr_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 ${FMAD:--fmad=false} -I. -cubin -o /tmp/one.cubin /root/repo/scratch/one.cu -Xptxas -v 2>&1 | grep -E "Used|error"
for k in 2 1 3; do echo "KIND $k"; cuobjdump -sass /tmp/one.cubin | awk -v k="Li${k}E" '/Function :/{f=($0 ~ k)} f' | grep -E "^\s+/\*[0-9a-f]+\*/" | awk '{print $2}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -18 | tr '\n' ' '; echo; done
