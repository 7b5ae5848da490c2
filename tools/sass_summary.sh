#!/bin/bash
# Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA use (B200_PROFILING.md), from the in-tree library.
# usage: tools/sass_summary.sh > profiles/rNN_sass_summary.txt
set -e
SO=${1:-tianshou_marl_b200/libtsm_b200.so}
echo "# cuobjdump -sass $SO  (arch: $(cuobjdump -lelf $SO | head -3 | tr '\n' ' '))"
echo "# kernel | UTCHMMA (tcgen05.mma) | LDTM (tcgen05.ld) | STTM (tcgen05.st) | UTCBAR (tcgen05.commit) | UTMALDG (TMA load) | UBLKCP (bulk copy) | SYNCS (mbarrier) | HMMA (legacy mma.sync)"
cuobjdump -sass "$SO" | awk '
/Function :/ { if (name != "") print name, u, l, s, b, t, k, y, h; name=$3; u=l=s=b=t=k=y=h=0 }
/UTCHMMA/ {u++} /LDTM/ {l++} /STTM/ {s++} /UTCBAR/ {b++} /UTMALDG/ {t++} /UBLKCP/ {k++} /SYNCS/ {y++} / HMMA/ {h++}
END { print name, u, l, s, b, t, k, y, h }' | while read n u l s b t k y h; do
  echo "$(echo $n | c++filt | cut -c1-90) | $u | $l | $s | $b | $t | $k | $y | $h"
done | sort -t'|' -k2 -n -r
