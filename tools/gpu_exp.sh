#!/bin/bash
cd /tmp
export STENCIL_LOG_LEVEL=1
A="/root/repo/build/bin/astaroth-sim --x 64 --y 48 --z 66 --fields 3 --double --star --iters 8 --check"
two() { # env assignments then args
  for r in 0 1; do RANK=$r WORLD_SIZE=2 LOCAL_RANK=$r MASTER_ADDR=127.0.0.1 MASTER_PORT=$PORT timeout 120 "$@" 2>/dev/null & done; wait
}
echo "1 proc free-run"; timeout 120 $A 2>/dev/null | tail -1
echo "1 proc host sync"; STENCIL_ASTAROTH_HOST_SYNC=1 timeout 120 $A 2>/dev/null | tail -1
echo "1 proc host sync no fused x"; STENCIL_ASTAROTH_HOST_SYNC=1 STENCIL_ASTAROTH_NO_FUSED_X=1 timeout 120 $A 2>/dev/null | tail -1
echo "1 proc 2 subdomains via staged only"; timeout 120 $A --remote 2>/dev/null | tail -1
echo "2 procs default"; PORT=31001 two $A
echo "2 procs no fused x"; export STENCIL_ASTAROTH_NO_FUSED_X=1; PORT=31101 two $A; unset STENCIL_ASTAROTH_NO_FUSED_X
echo "2 procs overlap schedule"; PORT=31201 two $A --overlap
echo "2 procs avg7"; PORT=31301 two /root/repo/build/bin/astaroth-sim --x 64 --y 48 --z 66 --fields 3 --double --iters 8 --check
echo "2 procs staged only"; PORT=31401 two $A --remote
