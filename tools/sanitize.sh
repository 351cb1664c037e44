#!/bin/bash
# Host code under AddressSanitizer + UndefinedBehaviorSanitizer (CPU only): the container code, encoders, writer and validator
# (csrc/host/*.cpp), the reader mirror (csrc/reader/*.cpp) and the oracle are rebuilt with -fsanitize=address,undefined into /tmp/fls_asan, linked with the
# already-built CUDA objects, and the CPU test suite (fuzzers included) runs against them.  Prints every sanitizer report;
# exit code 0 = the suite passed and nothing was reported.  The tree's own libraries are left alone.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=/tmp/fls_asan
mkdir -p $OUT/oracle
cd $ROOT/duckdb-fastlanes_b200/csrc
[ -d build ] || make -j8 > /dev/null
for f in host/*.cpp; do
	g++ -O1 -g -std=c++17 -fPIC -fsanitize=address,undefined -fno-omit-frame-pointer -ffp-contract=off -I../../include -Ihost -Ikernels \
		-c $f -o $OUT/$(basename $f .cpp).o &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libflsgpu.so $OUT/*.o build/*.cu.o -lcudart_static -lpthread -ldl -lrt \
	-Xlinker -lasan -Xlinker -lubsan
g++ -O1 -g -std=c++17 -fPIC -fsanitize=address,undefined -fno-omit-frame-pointer -ffp-contract=off -I../../include -Ihost -Ikernels -Ireader \
	-shared -o $OUT/libflsreader.so reader/*.cpp -L$OUT -lflsgpu -Wl,-rpath,'$ORIGIN' -lpthread
# the oracle is loaded from oracle/liboracle.so: run against a sanitized copy of the directory
cp $ROOT/oracle/*.c $ROOT/oracle/*.h $ROOT/oracle/Makefile $OUT/oracle/
gcc -O1 -g -fsanitize=address,undefined -fno-omit-frame-pointer -ffp-contract=off -fPIC -std=gnu11 -I$ROOT/include -shared \
	-o $OUT/oracle/liboracle.so $OUT/oracle/fls_oracle.c $OUT/oracle/fls_oracle_scan.c $OUT/oracle/fls_oracle_str.c -lpthread
cd $ROOT
export LD_PRELOAD="$(g++ -print-file-name=libasan.so) $(g++ -print-file-name=libubsan.so)"
export ASAN_OPTIONS=detect_leaks=0 FLSGPU_LIB=$OUT/libflsgpu.so FLS_ORACLE_LIB=$OUT/oracle/liboracle.so FLSREADER_LIB=$OUT/libflsreader.so
python -m pytest tests -q -s -m "not gpu" -p no:cacheprovider "$@" > $OUT/pytest.log 2>&1 || { tail -30 $OUT/pytest.log; exit 1; }
tail -1 $OUT/pytest.log
if grep -n "runtime error\|AddressSanitizer" $OUT/pytest.log; then exit 2; fi
echo "no sanitizer reports"
