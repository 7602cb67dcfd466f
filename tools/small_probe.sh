for rep in 1 2; do
for sel in "" "--only All"; do
  for nt in 100 300; do
    ./oracle/_ref/cuda-stream --csv --gigabytes -s 1048576 $sel -n $nt 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | tr '\n' ' ' | sed "s/^/[$sel n=$nt] ref        : /"; echo
    for m in stream auto; do
      BABELSTREAM_B200_SYNC=$m ./build/b200-stream --csv --gigabytes -s 1048576 $sel -n $nt 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | tr '\n' ' ' | sed "s/^/[$sel n=$nt] b200 $m : /"; echo
    done
  done
done
done
