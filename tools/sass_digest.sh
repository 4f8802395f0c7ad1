#!/bin/bash
# SASS evidence per tensor-core kernel of the shipped library: counts of the tcgen05 / TMEM / TMA mnemonics.
#   tools/sass_digest.sh > profiles/r2_sass_digest.md
so=voxelnet-pytorch_b200/lib/libvoxelnet_b200.so
echo "# SASS digest of $so ($(date -u +%F), nvcc $(nvcc --version | grep -o "release [0-9.]*"))"
echo
echo "\`cuobjdump -sass\` per kernel, \`grep -c\` of: UTCHMMA (tcgen05.mma), UTCBAR (tcgen05.commit), LDTM (tcgen05.ld), UTMALDG (TMA tensor load), UBLKCP (bulk copy), SYNCS (mbarrier), and registers / spill from \`cuobjdump -res-usage\`."
echo
echo "| kernel | UTCHMMA | UTCBAR | LDTM | UTMALDG | UBLKCP | SYNCS | regs | stack B |"
echo "|---|---|---|---|---|---|---|---|---|"
cuobjdump -sass $so > /tmp/_all.sass
cuobjdump -res-usage $so 2>/dev/null > /tmp/_res.txt
grep -n "Function : " /tmp/_all.sass | sed 's/[[:space:]]*Function : //' | while IFS=: read -r line fn; do
  echo "$line $fn"
done > /tmp/_fns.txt
total=$(wc -l < /tmp/_all.sass)
awk -v total=$total '{l[NR]=$1; f[NR]=$2} END {for (i=1;i<=NR;i++) print l[i], (i<NR? l[i+1]-1 : total), f[i]}' /tmp/_fns.txt | while read -r a b fn; do
  seg=$(sed -n "${a},${b}p" /tmp/_all.sass)
  mma=$(echo "$seg" | grep -c "UTCHMMA")
  ldtm=$(echo "$seg" | grep -c "LDTM")
  tma=$(echo "$seg" | grep -c "UTMALDG")
  if [ "$mma" -gt 0 ] || [ "$ldtm" -gt 0 ] || [ "$tma" -gt 0 ]; then
    bar=$(echo "$seg" | grep -c "UTCBAR"); blk=$(echo "$seg" | grep -c "UBLKCP"); syn=$(echo "$seg" | grep -c "SYNCS")
    res=$(grep -A1 "Function $fn:" /tmp/_res.txt | tail -1)
    regs=$(echo "$res" | grep -o "REG:[0-9]*" | cut -d: -f2); st=$(echo "$res" | grep -o "STACK:[0-9]*" | cut -d: -f2)
    echo "| \`$(echo $fn | c++filt | sed 's/vn::(anonymous namespace):://; s/(.*//')\` | $mma | $bar | $ldtm | $tma | $blk | $syn | $regs | $st |"
  fi
done
