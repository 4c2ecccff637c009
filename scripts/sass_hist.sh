#!/bin/bash
# static SASS opcode histogram of the production sweep kernels in an object file (development aid)
OBJ=${1:-lsc_internship_b200/_lib/obj/mhd_kernels.o}
for k in $(cuobjdump -elf $OBJ 2>/dev/null | grep -o "_ZN3lsc[0-9]*[a-z_]*sweep_kernelINS[A-Za-z0-9_]*" | grep -v _param_ | sort -u); do
  echo "== $k"
  cuobjdump -sass -fun $k $OBJ | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//; s/^@!?U?P[0-9T]+\s+//' \
   | awk '{split($1,a,"."); c[a[1]]++; n++} END{for(k in c) print c[k], k; print n, "TOTAL"}' | sort -rn | head -${2:-22} | tr '\n' ';'; echo
done
