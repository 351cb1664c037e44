for pf in 0 1; do for ctas in 1 2; do for ns in 1 2 3 4 6 8; do
echo -n "pf=$pf ctas=$ctas nslots=$ns : "; FLSGPU_PREFETCH=$pf FLSGPU_MAX_CTAS=$ctas FLSGPU_NSLOTS=$ns python tools/exp_cols.py 100000000 big,ts,int,date 2>&1 | grep "fwd"
done; done; done
