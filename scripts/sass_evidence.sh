#!/bin/bash
# SASS / PTX evidence for the Blackwell-native paths (B200_PROFILING.md "What proves a
# Blackwell-native kernel"): bulk async copies (UBLKCP), mbarrier (SYNCS), system-scope flags.
set -e
cd "$(dirname "$0")/.."
LIB=ddl_b200/lib/libddl_b200.so
OUT=profiles/sass_evidence.txt
{
  echo "# cuobjdump -sass $LIB  (built with -gencode arch=compute_100a,code=sm_100a -lineinfo)"
  cuobjdump -sass $LIB | awk '
    /Function :/ {f=$3}
    /UBLKCP/ {u[f]++} /SYNCS/ {s[f]++} /STG.E.STRONG.SYS|ST.E.STRONG.SYS/ {st[f]++}
    /LDG.E.STRONG.SYS|LD.E.STRONG.SYS/ {ld[f]++} /MEMBAR.*SYS/ {mb[f]++} /RED.E.*STRONG.SYS|REDG.*SYS/ {rd[f]++}
    /PRMT/ {p[f]++} /I2F/ {i2f[f]++} /MATCH/ {m[f]++} /LDG.E.*CONSTANT|LDG.E.128.CONSTANT/ {nc[f]++}
    END { printf "%-110s %7s %6s %8s %8s %7s %5s %5s %5s %6s\n","function","UBLKCP","SYNCS","ST.SYS","LD.SYS","MEMBAR","RED","PRMT","I2F","MATCH";
          for (k in u) seen[k]=1; for (k in s) seen[k]=1; for (k in st) seen[k]=1; for (k in ld) seen[k]=1; for (k in p) seen[k]=1; for (k in m) seen[k]=1;
          for (k in seen) printf "%-110s %7d %6d %8d %8d %7d %5d %5d %5d %6d\n", substr(k,1,110), u[k], s[k], st[k], ld[k], mb[k], rd[k], p[k], i2f[k], m[k] }' | sort
  echo
  echo "# PTX forms (cuobjdump -ptx): bulk copies, mbarrier, release/acquire at system scope"
  cuobjdump -ptx $LIB | grep -oE "cp\.async\.bulk[a-z0-9_.:]*|mbarrier\.[a-z0-9_.:]*|st\.release\.sys[a-z0-9_.]*|ld\.acquire\.sys[a-z0-9_.]*|red\.release\.sys[a-z0-9_.]*|fence\.[a-z0-9_.:]*|multimem\.[a-z0-9_.]*" | sort | uniq -c | sort -rn
  echo
  echo "# one UBLKCP site in context (gather_tma_copy_kernel)"
  cuobjdump -sass $LIB | grep -B2 -A2 "UBLKCP" | head -24
} > $OUT
echo wrote $OUT
