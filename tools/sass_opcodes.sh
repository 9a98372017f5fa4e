#!/bin/bash
# tools/sass_opcodes.sh > profiles/r2_sass_opcodes.txt : opcode histogram of the shipped library (cuobjdump -sass), Blackwell-only opcodes first
LIB=${1:-vitvs_b200/libvitvs_b200.so}
echo "# $(basename $LIB) build id $(python -c 'from vitvs_b200 import build; print(build.binary_id())'), $(cuobjdump -lelf $LIB | grep -c sm_100a) sm_100a cubins; nvcc $(nvcc --version | grep -o 'release [0-9.]*')"
cuobjdump -sass $LIB | awk '$1 ~ /^\/\*[0-9a-f]+\*\/$/ {op=$2; if (op ~ /^@/) op=$3; sub(/;/,"",op); print op}' > /tmp/ops.txt
echo "# tcgen05 / TMEM / TMA / cluster / mbarrier opcodes (full mnemonic):"
grep -E "^(UTC|UTMA|LDTM|STTM|UTCBAR|SYNCS|UCGABAR|REDUX|MUFU|UBLKCP|UBLKRED|CCTL|FENCE|ACQBULK|ARRIVES|ELECT|UGETNEXTWORKID|USETMAXREG)" /tmp/ops.txt | sort | uniq -c | sort -rn
echo "# all opcodes (base mnemonic):"
sed 's/\..*//' /tmp/ops.txt | sort | uniq -c | sort -rn
