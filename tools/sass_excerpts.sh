#!/bin/bash
# Writes profiles/r02_sass_<kernel>.txt: per hot kernel the resource line and the SASS lines that show the
# mechanisms DESIGN.md names (bulk copies + mbarrier, LDGSTS staging, programmatic dependent launch, FP64 tensor-core MMA).
set -e
LIB=dae-cpp_b200/libdaecpp_b200.so
OUT=profiles
dump() { # $1 = mangled-name regex, $2 = file tag, $3 = grep pattern
  local fn
  fn=$(cuobjdump -sass $LIB | grep -E "Function : .*$1" | head -1 | sed 's/.*Function : //')
  {
    echo "# $fn"
    echo "# cuobjdump -sass $LIB  (sm_100a), lines matching: $3"
    cuobjdump -res-usage $LIB 2>/dev/null | grep -A1 -F "$fn" | tail -1
    cuobjdump -sass -fun "$fn" $LIB 2>/dev/null | grep -nE "$3" | head -${4:-60}
    echo "# instruction counts:"
    cuobjdump -sass -fun "$fn" $LIB 2>/dev/null | grep -oE "^\s+/\*[0-9a-f]+\*/\s+[A-Z0-9_.]+" | awk '{print $2}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -25
  } > $OUT/r02_sass_$2.txt
}
dump "k_chain_factorILi1ELb1" chain_factor_b1 "UBLKCP|SYNCS|LDGSTS|ACQBULK|PREEXIT|DEPBAR|MUFU.RCP64H|BAR.SYNC" 80
dump "k_chain_factorILi2ELb1" chain_factor_b2 "UBLKCP|SYNCS|LDGSTS|ACQBULK|PREEXIT|DEPBAR|BAR.SYNC" 60
dump "k_chain_forwardILi1ELb1" chain_forward_b1 "LDGSTS|ACQBULK|PREEXIT|DEPBAR|BAR.SYNC|CCTL|LDG.E.*128" 60
dump "k_chain_backwardILi1ELb1" chain_backward_b1 "LDGSTS|ACQBULK|PREEXIT|DEPBAR|BAR.SYNC|CCTL|STG.E|MEMBAR" 60
dump "k_bf_gemmILb1" bf_gemm_dmma "DMMA|ACQBULK|PREEXIT|LDS|BAR.SYNC" 60
dump "k_unit_factor" unit_factor "DMMA|DFMA|BAR.SYNC|ATOM|RED|LDGSTS" 40
