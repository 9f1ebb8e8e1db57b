#!/bin/bash
# Per-kernel histogram of the SASS opcodes that prove a Blackwell-native kernel (B200_PROFILING.md): tcgen05.mma ->
# UTC*MMA, tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG/UTMASTG/UBLKCP; legacy tensor path HMMA must be absent.
#   bash profiles/sass_histogram.sh > profiles/r02_sass_histogram.txt
SO=${1:-dphubert_b200/libdphubert_b200.so}
cuobjdump -sass $SO | awk '
/Function :/ { f=$3; sub(/^_ZN[0-9]*[a-z0-9]*[0-9]+/, "", f); fn=$3 }
/UTC[A-Z]*MMA/ { mma[fn]++ } /LDTM/ { ldtm[fn]++ } /STTM/ { sttm[fn]++ } /UTMALDG/ { tld[fn]++ } /UTMASTG/ { tst[fn]++ }
/UBLKCP/ { blk[fn]++ } / HMMA/ { hmma[fn]++ } /Function :/ { all[fn]=1 }
END { for (k in all) if (mma[k]+ldtm[k]+tld[k]+tst[k]+blk[k]+hmma[k] > 0)
        printf "%5d UTC*MMA %4d LDTM %4d STTM %4d UTMALDG %4d UTMASTG %3d UBLKCP %3d HMMA  %s\n", mma[k], ldtm[k], sttm[k], tld[k], tst[k], blk[k], hmma[k], k }' | sort -k9 | c++filt -p 2>/dev/null | sed 's/  */ /g' | cut -c1-230
echo "# totals:"; cuobjdump -sass $SO | grep -oE "UTC[A-Z]*MMA|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP| HMMA" | sort | uniq -c
