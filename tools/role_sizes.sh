#!/bin/bash
# SASS size of each growth role of planar_grow_kernel<384,true,false> (roles made noinline in a scratch copy)
set -e
T=/tmp/pcseg_sizes; rm -rf $T; mkdir -p $T; cp -r /root/repo/x-edr-pointcloud-segmentation_b200/csrc $T/
cd $T/csrc
sed -i 's|#include "../../include/pcseg_b200.h"|#include "/root/repo/include/pcseg_b200.h"|' pcseg_api.cu
sed -i 's/__device__ __forceinline__ void pipe_warp_A/__device__ __noinline__ void pipe_warp_A/; s/__device__ __forceinline__ void pipe_warp_S/__device__ __noinline__ void pipe_warp_S/; s/__device__ __forceinline__ void pipe_warp_R/__device__ __noinline__ void pipe_warp_R/; s/__device__ __forceinline__ void pipe_group_B/__device__ __noinline__ void pipe_group_B/' grow_planar_pipe.cuh
sed -i 's/__device__ __forceinline__ RawFit est_raw/__device__ __noinline__ RawFit est_raw/' pcseg_math.cuh
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -cubin -o $T/ni.cubin pcseg_api.cu
nvdisasm $T/ni.cubin > $T/ni.sass
python - <<'PY'
import re
sec=None; lab=None; sizes={}
for line in open('/tmp/pcseg_sizes/ni.sass'):
    m=re.match(r'\.text\.(\S+):',line)
    if m: sec=m.group(1); lab='<main>'; continue
    m=re.match(r'(\$?_Z\S+|\$\S+):',line.strip())
    if m and sec:
        if not m.group(1).startswith('.L'): lab=m.group(1)
        continue
    if sec and ('planar_grow_kernelILi' in sec and 'ELb1ELb0' in sec) and re.match(r'\s+/\*[0-9a-f]{4,}\*/',line):
        sizes[lab]=sizes.get(lab,0)+16
for k,v in sorted(sizes.items(), key=lambda x:-x[1])[:9]: print("%6.1f KB  %s"%(v/1024,k.split('$')[-1][:60]))
PY
