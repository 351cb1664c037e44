for r in 1 2 3; do
  for lib in libflsgpu.so libflsgpu_plain.so; do
    FLSGPU_LIB=$PWD/duckdb-fastlanes_b200/$lib python tools/ab_scan.py 100000000 30 2>&1 | tail -1
  done
done
