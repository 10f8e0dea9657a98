#!/bin/bash
mkdir -p gpurun_out; exec 2>&1
L=nbody_b200/libnbody_b200
bash tools/ab_direct.sh $L.so ${L}_occ6.so ${L}_i4occ4.so ${L}_i4occ5.so ${L}_i4t128.so ${L}_i4occ4u32.so ${L}_occ6u32.so ${L}_occ6u128.so ${L}_occ6jt512.so ${L}_occ6.so > gpurun_out/ab_direct2.log 2>&1
AB_MIXED=1 bash tools/ab_direct.sh $L.so ${L}_prev.so ${L}_occ6.so ${L}_i4occ4.so ${L}_occ4.so >> gpurun_out/ab_direct2.log 2>&1
cat gpurun_out/ab_direct2.log
