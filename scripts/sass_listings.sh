#!/bin/bash
# Full SASS listings of the main kernels into profiles/sass/ (cuobjdump -sass -fun <mangled name>).
cd "$(dirname "$0")/.."
LIB=ddl_b200/lib/libddl_b200.so
OUT=profiles/sass
mkdir -p $OUT
names=$(cuobjdump -sass $LIB | grep "Function :" | awk '{print $3}')
pick() { echo "$names" | grep -E "$1" | head -1; }
dump() { local f; f=$(pick "$1"); [ -n "$f" ] && cuobjdump -sass -fun "$f" $LIB 2>/dev/null | grep -v "^\s*/\* 0x" > $OUT/$2.sass && echo "$2 <- $f"; }
dump 'gather_tma_cvt_kernelILi1ELb1ELi4096ELi3ELi16E' gather_tma_cvt_u8_bf16_chan_4096x3x16
dump 'gather_tma_copy_kernelILi8192ELi3E' gather_tma_copy_8192x3
dump 'gather_ldg_kernelILi1ELb1E' gather_ldg_u8_bf16_chan
dump 'push_tma_cvt_kernelILi1ELb1ELi4096ELi3ELi16ELb0E' push_tma_cvt_u8_bf16_chan_unicast
dump 'push_tma_cvt_kernelILi1ELb1ELi4096ELi3ELi16ELb1E' push_tma_cvt_u8_bf16_chan_multicast
dump 'perm_gen_kernel' perm_gen
dump 'index_exchange_kernel' index_exchange
dump 'flag_signal_kernel' flag_signal
dump 'flag_wait_kernel' flag_wait
dump 'flag_barrier_kernel' flag_barrier
dump 'multicast_put_kernel' multicast_put
dump 'checksum_kernel' checksum
