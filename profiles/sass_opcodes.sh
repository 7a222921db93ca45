#!/bin/bash
# per-kernel counts of the SASS mnemonics that prove the Blackwell path (B200_PROFILING.md "What proves a Blackwell-native kernel")
# usage: profiles/sass_opcodes.sh > profiles/r02_sass_opcodes.txt
LIB=${1:-gnn_manip_b200/libgnn_manip_b200.so}
echo "# cuobjdump -sass $LIB : UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk (TMA unit, 1-D), UTCBAR = tcgen05.commit, LDGSTS = cp.async, HMMA = legacy mma.sync (must be 0)"
cuobjdump -sass "$LIB" | awk '
/Function :/ { fn=$3; sub(/^_ZN2gm[0-9]*_GLOBAL__N__[0-9a-f_]*cu_[0-9a-f]*/, "", fn); names[fn]=1 }
{ for (i=1;i<=NF;i++) { op=$i; sub(/\..*/, "", op);
    if (op ~ /^UTC[A-Z]*MMA$/) c[fn,"UTCMMA"]++;
    else if (op=="LDTM"||op=="STTM"||op=="UBLKCP"||op=="UTCBAR"||op=="LDGSTS"||op=="HMMA"||op=="UTMALDG"||op=="UTMASTG"||op=="RED"||op=="REDG") c[fn,op]++ } }
END { printf "%-30s %7s %5s %5s %6s %6s %6s %5s\n","kernel","UTC*MMA","LDTM","STTM","UBLKCP","UTCBAR","LDGSTS","HMMA";
  for (f in names) if (c[f,"UTCMMA"]+c[f,"LDTM"]+c[f,"UBLKCP"]+c[f,"LDGSTS"]+c[f,"HMMA"]>0)
    printf "%-30s %7d %5d %5d %6d %6d %6d %5d\n", f, c[f,"UTCMMA"],c[f,"LDTM"],c[f,"STTM"],c[f,"UBLKCP"],c[f,"UTCBAR"],c[f,"LDGSTS"],c[f,"HMMA"] }' | sort
