#!/bin/bash
# SASS mnemonic census of the bulk kernels (cuobjdump on the built library): what proves the data movement and the atomics
# each kernel uses.  usage: tools/sass_evidence.sh > profiles/rNN_sass_bulk_kernels.txt
LIB=kmerhash_b200/libkmerhash_b200.so
for pat in 'l1_scatter_kernelILi1E' 'l1_scatter_kernelILi2E' 'l2_scatter_kernelILi1E' 'tile_insert_kernelILi1E' 'tile_insert_kernelILi2E' \
           'tile_rebuild_kernelILi1E' 'kmer_tiles_kernelILi1ELb0E' 'kmer_tiles_kernelILi1ELb1E' 'scatter_slabs_kernelILi1ELb0ELb0E' 'peer_sync_kernel'; do
  fn=$(cuobjdump -sass $LIB 2>/dev/null | grep -E "Function : .*${pat}" | head -n 1 | sed 's/.*Function : //')
  [ -z "$fn" ] && continue
  echo "== $fn"
  cuobjdump -sass -fun "$fn" $LIB 2>/dev/null | grep -E "^\s+/\*[0-9a-f]{4,}\*/" | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//' | sed -E 's/^@!?U?P[0-9T]+ //' | awk '{print $1}' | sed 's/;//' \
    | grep -E "^(UBLKCP|UTMA|SYNCS|ATOMS|ATOMG|ATOM|RED|LDG|STG|LDS|STS|BAR|MATCH|VOTE|REDUX|SHFL|MEMBAR|FENCE|IMAD\.WIDE|BREV|POPC|FLO|NANOSLEEP)" | sort | uniq -c | sort -rn | awk '{printf "   %6d %s\n", $1, $2}'
  echo "   registers/spills: see profiles/*ncu* (launch__registers_per_thread) and nvcc -Xptxas -v"
done
