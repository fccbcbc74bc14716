#!/bin/bash
# Per-kernel counts of the Blackwell-native SASS mnemonics (tcgen05 MMA / TMEM load-store / TMA load-store / commit barriers)
# in the built library.  Usage: bash tools/sass_counts.sh > profiles/r02/sass_counts.txt
SO=${1:-mdgen_finetune_b200/libmdgen_b200.so}
echo "cuobjdump -sass $SO (sm_100a cubin): instruction counts per kernel"
printf "%-72s %8s %6s %6s %8s %8s %7s\n" kernel UTCHMMA LDTM STTM UTMALDG UTMASTG UTCBAR
cuobjdump -sass "$SO" | awk '
  /Function :/ { if (name != "") out(); name=$3; a=b=c=d=e=f=0 }
  /UTCHMMA/ {a++} /LDTM/ {b++} /STTM/ {c++} /UTMALDG/ {d++} /UTMASTG/ {e++} /UTCBAR/ {f++}
  function out() { if (a+b+c+d+e+f > 0) printf "%-72s %8d %6d %6d %8d %8d %7d\n", substr(name,1,72), a,b,c,d,e,f; ta+=a;tb+=b;tc+=c;td+=d;te+=e;tf+=f }
  END { out(); printf "%-72s %8d %6d %6d %8d %8d %7d\n", "TOTAL", ta,tb,tc,td,te,tf }' | (command -v c++filt >/dev/null && c++filt || cat)
