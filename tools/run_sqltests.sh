#!/bin/bash
# Runs every SQLLogicTest of the extension one file at a time under its own timeout (a hung scan must not eat the GPU box's
# whole time limit); logs under gpurun_out/.
cd "$(dirname "$0")/.."
export LD_LIBRARY_PATH=$PWD/duckdb-fastlanes_b200/extension/bin:$PWD/duckdb-fastlanes_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
for f in duckdb-fastlanes_b200/extension/test/sql/*.test; do
  n=$(basename $f .test)
  rm -rf /tmp/fls_sqltest_$n
  timeout ${SQLTEST_TIMEOUT:-180} duckdb-fastlanes_b200/extension/bin/unittest --test-dir duckdb-fastlanes_b200/extension \
    --test-temp-dir /tmp/fls_sqltest_$n "test/sql/$n.test" > gpurun_out/sqltest_$n.log 2>&1
  echo "$n: exit $?"; tail -25 gpurun_out/sqltest_$n.log
done
