#!/bin/bash
# profiles/r2_sass_<object>.txt: per kernel, how often each memory-system / synchronisation
# mnemonic occurs in the SASS of the built objects (needs no GPU: cuobjdump reads the .o files).
#   UBLKCP.S.G / .G.S = cp.async.bulk (TMA), UBLKPF.L2 = bulk L2 prefetch, SYNCS.* = mbarrier,
#   LDGSTS = cp.async, STTM / LDTM = tcgen05.st / tcgen05.ld (tensor memory), UTCATOMSWS = TMEM alloc
cd "$(dirname "$0")/../nnstreamer_b200/csrc/build" || exit 1
for o in k_stand_fused k_plane k_perm_vec k_perm_slab k_perm_pix k_perm k_chain_row_5 k_clamp k_stand_row_7; do
  cuobjdump -sass $o.o | grep -E "Function :|UBLKCP|UBLKPF|SYNCS|LDGSTS|LDG\.E.*128|LDG\.E.*256|STG\.E.*128|STG\.E.*256|STTM|LDTM|UTCATOMSWS|LDS\.128|STS\.128|ATOMG|REDG" \
  | awk '/Function :/{fn=$0; next} {gsub(/\/\*[0-9a-f]+\*\//,""); sub(/^[ \t]+/,""); split($0,a," "); op=a[1]; if (op ~ /^@/) op=a[2]; cnt[fn"|"op]++} END{for (k in cnt){split(k,b,"|"); print b[1] "  " b[2] " x" cnt[k]}}' \
  | sed 's/^[ \t]*Function : //' | sort > ../../../profiles/r2_sass_$o.txt
done
