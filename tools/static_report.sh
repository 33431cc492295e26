#!/bin/bash
# Static evidence that needs no GPU: ptxas resource usage per kernel (registers, shared memory,
# spills) and the SASS mnemonics that prove which data path each kernel uses (UBLKCP = 1-D bulk
# async copy, SYNCS = mbarrier traffic, LDG/STG widths). Usage: tools/static_report.sh > profiles/<name>.txt
set -e
cd "$(dirname "$0")/../rela_b200/csrc"
TMP=$(mktemp -d)
echo "# nvcc $(nvcc --version | grep release | sed 's/.*release //') -gencode arch=compute_100a,code=sm_100a -lineinfo -O3"
echo
echo "## ptxas -v"
for f in row_copy select gather_convert actor_math; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xptxas -v -c $f.cu -o $TMP/$f.o 2>&1 |
    awk '/Compiling entry function/ {name=$0; sub(/.*function .?/, "", name); sub(/. for .*/, "", name)}
         /bytes stack frame/ {st=$0; sub(/^ +/, "", st)}
         /Used [0-9]+ registers/ {u=$0; sub(/.*: /, "", u); printf "%s\n    %s | %s\n", name, u, st}' | c++filt
done
echo
echo "## SASS mnemonic counts per kernel (cuobjdump -sass of the object files)"
for f in row_copy select gather_convert actor_math; do
  cuobjdump -sass $TMP/$f.o | awk '
    /Function :/ {fn=$3}
    /^ +\/\*[0-9a-f]+\*\/ +/ {
      line=$0; sub(/^ +\/\*[0-9a-f]+\*\/ +/, "", line); sub(/^@!?U?P[0-9T]+ +/, "", line);
      split(line, a, /[ ;]/); op=a[1];
      if (op ~ /^(UBLKCP|SYNCS|LDG|STG|LDS|STS|ATOM|ATOMG|RED|ACQBULK|UTMA|DADD|DMUL|DFMA|SHFL|BAR|ERRBAR|MEMBAR|CCTL|LDGDEPBAR|DEPBAR|WARPSYNC|VOTE|REDUX)/) c[fn" "op]++
    }
    END {for (k in c) print c[k], k}' | sort -k2,2 -k3,3 | awk '{printf "%-8s %s  %s\n", $1, $3, $2}' | c++filt
done
rm -rf $TMP
