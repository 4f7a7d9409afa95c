#!/bin/sh
# ORACLE BUILD ONLY. usage: mk_blas_alias.sh <archive> <blas.so> <prefix> > alias.S
# Newer OpenBLAS wheels export the Fortran entry points with a prefix (scipy_dgemm_); the
# reference calls the plain names (dgemm_).  Emit one tail-jump per Fortran symbol the archive
# leaves undefined and the library exports under the prefix — arguments pass through untouched.
ar=$1; so=$2; pre=$3
nm -D --defined-only "$so" | awk '{print $3}' | grep "^${pre}" | sed "s/^${pre}//" | sort -u > /tmp/_elb_blas_exports.$$
nm -u "$ar" 2>/dev/null | awk '$1=="U"{print $2}' | grep -E '^[a-z0-9_]+_$' | sort -u > /tmp/_elb_blas_needed.$$
echo '.text'
comm -12 /tmp/_elb_blas_exports.$$ /tmp/_elb_blas_needed.$$ | while read s; do
  printf '.globl %s\n.type %s,@function\n%s:\n\tjmp %s%s@PLT\n' "$s" "$s" "$s" "$pre" "$s"
done
echo '.section .note.GNU-stack,"",@progbits'
rm -f /tmp/_elb_blas_exports.$$ /tmp/_elb_blas_needed.$$
