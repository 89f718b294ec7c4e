#!/bin/bash
# Counts the tensor-core / TMA / async-copy SASS opcodes per kernel of libwdmpnn.so (profiles/sass_opcodes.txt).
lib=${1:-polymer_chemprop_b200/libwdmpnn.so}
echo "# cuobjdump -sass $lib (sm_100a): opcode counts per kernel"
echo "# UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTMALDG = TMA tensor load, UBLKCP = bulk copy,"
echo "# LDGSTS = cp.async, SYNCS = mbarrier ops, BRA.U.ANY = per-MMA waterfall loops (0 on the hot MMA paths since r2)"
cuobjdump -sass "$lib" | awk '
  /Function :/ { name=$3 }
  { for (k in ops) if (index($0, k)) cnt[name, k]++ }
  BEGIN { split("UTCHMMA LDTM UTCBAR UTMALDG UBLKCP LDGSTS SYNCS BRA.U.ANY", o, " "); for (i in o) ops[o[i]]=1 }
  /Function :/ { names[name]=1 }
  END { for (n in names) { line=""; tot=0; for (i=1;i<=8;i++) { k=o[i]; c=cnt[n,k]+0; if (k!="SYNCS" && k!="BRA.U.ANY") tot+=c; line=line sprintf(" %s=%d", k, c) } if (tot>0) print n ":" line } }' | c++filt | sort
