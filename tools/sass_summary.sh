#!/bin/sh
# Per-kernel counts of the SASS mnemonics that prove a Blackwell-native kernel (tcgen05.mma -> UTC*MMA, tcgen05.ld/st ->
# LDTM / STTM, TMA -> UTMALDG / UBLKCP / UBLKRED) in the shipped library.  Usage: tools/sass_summary.sh > profiles/rN_sass_summary.txt
LIB=${1:-orange3-spark_b200/b200fit/libb200fit.so}
echo "# cuobjdump -sass $LIB, per kernel: mnemonic counts (legacy HMMA / HGMMA would show a non-native tensor path)"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { name=$3 }
  /UTCHMMA|UTCQMMA|UTCIMMA|UTCOMMA/ { mma[name]++ }
  /LDTM/ { ldtm[name]++ }
  /STTM/ { sttm[name]++ }
  /UTMALDG/ { tma[name]++ }
  /UBLKCP/ { blk[name]++ }
  /UBLKRED|UTMAREDG/ { red[name]++ }
  /HMMA|HGMMA/ { if ($0 !~ /UTC/) legacy[name]++ }
  /Function :/ { seen[name]=1 }
  END {
    printf "%-10s %-6s %-6s %-8s %-7s %-8s %-7s %s\n", "UTC*MMA", "LDTM", "STTM", "UTMALDG", "UBLKCP", "UBLKRED", "legacy", "kernel"
    for (k in seen) if (mma[k] + ldtm[k] + sttm[k] + tma[k] + blk[k] + red[k] + legacy[k] > 0)
      printf "%-10d %-6d %-6d %-8d %-7d %-8d %-7d %s\n", mma[k], ldtm[k], sttm[k], tma[k], blk[k], red[k], legacy[k], k
  }' | (read hdr; echo "$hdr"; sort -k8)
