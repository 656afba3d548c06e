#!/bin/bash
# static SASS summary of one kernel of a library: instruction count, local-memory accesses, FP64 ops, stores by width
# usage: sasscount.sh lib.so 'rollout_kernelILi3ELb1'
LIB=$1; K=$2
cuobjdump -sass -fun "$(cuobjdump -elf $LIB 2>/dev/null | grep -o "_Z[A-Za-z0-9_]*${K}[A-Za-z0-9_]*" | sort -u | head -1)" $LIB 2>/dev/null > /tmp/sass_$$.txt
awk '/^\s+\/\*[0-9a-f]{4}\*\//{n++} END{print "instructions:", n}' /tmp/sass_$$.txt
for pat in LDL STL "LDG\|LD\.E" "STG\|ST\.E" LDS STS ATOMS "DADD\|DMUL\|DFMA\|DSETP\|DMNMX" "IMAD.MOV" SHFL "BSSY\|BSYNC" "BRA" MUFU "F2F\|I2F\|F2I\|FRND"; do
  printf "%-40s %s\n" "$pat" "$(grep -c "$pat" /tmp/sass_$$.txt)"
done
rm -f /tmp/sass_$$.txt
