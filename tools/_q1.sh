export LD_LIBRARY_PATH=$PWD/duckdb-fastlanes_b200/extension/bin:$PWD/duckdb-fastlanes_b200
D=duckdb-fastlanes_b200/extension/bin/duckdb
rm -f /dev/shm/q1.duckdb /dev/shm/q1.fls
$D /dev/shm/q1.duckdb -c "CALL dbgen(sf=10); COPY lineitem TO '/dev/shm/q1.fls' (FORMAT fls);" > /dev/null
Q1="SELECT l_returnflag, l_linestatus, sum(l_quantity), sum(l_extendedprice), sum(l_extendedprice * (1 - l_discount)), sum(l_extendedprice * (1 - l_discount) * (1 + l_tax)), avg(l_quantity), avg(l_extendedprice), avg(l_discount), count(*) FROM SRC WHERE l_shipdate <= CAST('1998-09-02' AS date) GROUP BY l_returnflag, l_linestatus ORDER BY l_returnflag, l_linestatus"
for src in "lineitem" "read_fls('/dev/shm/q1.fls')"; do
  q=${Q1/SRC/$src}
  $D /dev/shm/q1.duckdb -c ".timer on" -c "$q" -c "$q" -c "EXPLAIN ANALYZE $q" 2>&1 | grep -E "Run Time|Total Time|READ_FLS|SEQ_SCAN|HASH_GROUP|PROJECTION|\(.*s\)" | head -30
  echo ----
done
for t in 8 32; do for b in 4 8 16; do
  q=${Q1/SRC/"read_fls('/dev/shm/q1.fls')"}
  echo "threads $t batch $b"; $D /dev/shm/q1.duckdb -c "SET threads=$t; SET fastlanes_rowgroups_per_batch=$b;" -c ".timer on" -c "$q" -c "$q" -c "$q" 2>&1 | grep "Run Time" | tail -2
done; done
q=${Q1/SRC/"read_fls('/dev/shm/q1.fls')"}
echo "select_on_device off"; $D /dev/shm/q1.duckdb -c "SET fastlanes_select_on_device=false;" -c ".timer on" -c "$q" -c "$q" -c "$q" 2>&1 | grep "Run Time" | tail -2
rm -f /dev/shm/q1.duckdb /dev/shm/q1.fls
