#!/bin/bash
# usage: build_variant.sh NAME "-DFLAGS"
set -e
cd /root/repo/distributed-gpu-test_b200/csrc
mkdir -p /root/repo/tools/bin/var_$1
for f in *.cu; do
  if [ "$f" = "mlp_tc.cu" ]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr $2 -c $f -o /root/repo/tools/bin/var_$1/${f%.cu}.o &
  fi
done
wait
OBJS=""
for f in *.cu; do
  if [ "$f" = "mlp_tc.cu" ]; then OBJS="$OBJS /root/repo/tools/bin/var_$1/mlp_tc.o"; else OBJS="$OBJS build/${f%.cu}.o"; fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o /root/repo/tools/bin/lib_$1.so $OBJS -ldl
echo built lib_$1.so
