#!/bin/bash
# Per kernel of libv2m_cuda.so: the SASS mnemonics that show what the kernel is built on (B200_PROFILING.md: UBLKCP = TMA 1-D
# bulk copies, SYNCS = mbarrier operations), plus the instruction count. Usage: scripts/sass_evidence.sh > profiles/rNN_sass.txt
LIB=${1:-vcf2multialign_b200/libv2m_cuda.so}
echo "# cuobjdump -sass $LIB ($(cuobjdump -lelf $LIB | grep -c cubin) cubin(s): $(cuobjdump -lelf $LIB | grep -o 'sm_[0-9a-z]*' | sort -u | tr '\n' ' '))"
echo "# instructions | UBLKCP.S.G (bulk global->shared) | UBLKCP.G.S (bulk shared->global) | SYNCS.* (mbarrier) | REDUX/CREDUX | SHFL | VOTE | HSET2 | VIMNMX3 | LDG | STG | LDS | STS | kernel"
cuobjdump -sass $LIB | awk '
/Function : / { if (fn != "") flush(); fn=$3; delete c; n=0 }
/^ +\/\*[0-9a-f]{4}\*\// { n++; for (i=2;i<=NF;i++) if ($i !~ /^@/ && $i !~ /^\/\*/) { op=$i; break }
  if (op ~ /^UBLKCP\.S\.G/) c["sg"]++; else if (op ~ /^UBLKCP\.G\.S/) c["gs"]++; else if (op ~ /^SYNCS/) c["sy"]++;
  else if (op ~ /REDUX/) c["rx"]++; else if (op ~ /^SHFL/) c["sh"]++; else if (op ~ /^VOTE/) c["vo"]++; else if (op ~ /^HSET2/) c["hs"]++;
  else if (op ~ /^VIMNMX3/) c["vm"]++; else if (op ~ /^LDG/) c["ldg"]++; else if (op ~ /^STG/) c["stg"]++; else if (op ~ /^LDS/) c["lds"]++; else if (op ~ /^STS/) c["sts"]++ }
function flush() { printf "%6d | %3d | %3d | %4d | %3d | %4d | %3d | %3d | %3d | %4d | %4d | %4d | %4d | %s\n", n, c["sg"], c["gs"], c["sy"], c["rx"], c["sh"], c["vo"], c["hs"], c["vm"], c["ldg"], c["stg"], c["lds"], c["sts"], fn }
END { flush() }' | c++filt | sed 's/(.*//' | sort -t'|' -k14
