#!/bin/bash
# Appends the instruction evidence of the hot kernels to profiles/r02_sass_excerpt.md
SO=kube_gpu_stats_b200/libkgs.so
OUT=profiles/r02_sass_excerpt.md
fn() { cuobjdump -sass $SO 2>/dev/null | awk -v k="$1" '/Function : /{f=index($0,k)>0} f'; }
{
echo; echo '## `k_accum<PCT=1, TMA=0>`: 256-bit streaming loads of the four sample columns'; echo '```'
fn k_accumILb1ELb0 | grep -E "LDG\.E.*256|arch =|Function" | head -8; echo '```'
echo; echo '## `k_accum<PCT=1, TMA=1>` (optional path): bulk copies + mbarrier'; echo '```'
fn k_accumILb1ELb1 | grep -E "UBLKCP|SYNCS" | head -8; echo '```'
echo; echo '## `k_finalize_groups<8>`: dp4a rank counting (IDP.4A.S8.U8), group sums by SHFL.BFLY, PDL (ACQBULK / griddepcontrol)'; echo '```'
fn k_finalize_groupsILi8 | grep -E "IDP\.4A" | head -4
fn k_finalize_groupsILi8 | grep -E "SHFL\.BFLY" | head -3
fn k_finalize_groupsILi8 | grep -E "ACQBULK|PREEXIT|DEPBAR|griddep" | head -4
echo "IDP.4A count: $(fn k_finalize_groupsILi8 | grep -c 'IDP\.4A')  total instructions: $(fn k_finalize_groupsILi8 | grep -c '/\*[0-9a-f]\{4\}\*/')"; echo '```'
echo; echo '## `k_peer_barrier`: system-scope release/acquire on peer-mapped flag words'; echo '```'
fn k_peer_barrier | grep -E "\.SYS|MEMBAR|NANOSLEEP" | head -8; echo '```'
echo; echo '## registers / shared memory (`cuobjdump -res-usage`)'; echo '```'
cuobjdump -res-usage $SO 2>/dev/null | grep -A1 -E "k_accumILb1ELb0|k_finalize_groupsILi8|k_finalize_cells|k_finalize_rare|k_finalize_slide|k_peer_barrier" | grep -E "Function|REG" | sed 's/^ *//' | cut -c1-150; echo '```'
} >> $OUT
