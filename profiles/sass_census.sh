#!/bin/bash
# SASS census of the built library: Blackwell-specific instruction families per kernel.
#   bash profiles/sass_census.sh > profiles/r02_sass_census.txt
SO=${1:-udp_pose_b200/libudp_codec.so}
cuobjdump -sass $SO > /tmp/udp_sass.txt
echo "library: $SO   $(grep -c 'Function :' /tmp/udp_sass.txt) kernels   archs: $(grep "arch =" /tmp/udp_sass.txt | sort | uniq -c | tr -s " " | tr "\n" ";")"
echo "totals: UTMALDG $(grep -c UTMALDG /tmp/udp_sass.txt)  SYNCS $(grep -c 'SYNCS\.' /tmp/udp_sass.txt)  FFMA2 $(grep -c FFMA2 /tmp/udp_sass.txt)  FADD2 $(grep -c FADD2 /tmp/udp_sass.txt)  FMUL2 $(grep -c FMUL2 /tmp/udp_sass.txt)  FMNMX3 $(grep -c FMNMX3 /tmp/udp_sass.txt)  IDP.2A $(grep -c 'IDP\.2A' /tmp/udp_sass.txt)  DFMA $(grep -c DFMA /tmp/udp_sass.txt)  UTC*MMA $(grep -c 'UTC.*MMA' /tmp/udp_sass.txt) (no contraction on this path)"
echo
printf "%-110s %7s %8s %6s %6s %6s %7s %7s %6s\n" kernel instrs UTMALDG SYNCS FFMA2 FADD2 FMNMX3 IDP.2A DFMA
awk '
/Function :/ { if (name != "") printf "%-110s %7d %8d %6d %6d %6d %7d %7d %6d\n", substr(name,1,110), n, t, s, f, a, m, d, g;
               name=$3; n=t=s=f=a=m=d=g=0; next }
/^ +\/\*[0-9a-f]+\*\/ / { n++ }
/UTMALDG/ {t++} /SYNCS\./ {s++} /FFMA2/ {f++} /FADD2/ {a++} /FMNMX3/ {m++} /IDP\.2A/ {d++} /DFMA/ {g++}
END { printf "%-110s %7d %8d %6d %6d %6d %7d %7d %6d\n", substr(name,1,110), n, t, s, f, a, m, d, g }' /tmp/udp_sass.txt | c++filt -p 2>/dev/null | sort
