#!/bin/bash
# end-of-round verification of the committed state: GPU tests, smoke(), the default bench line, the reference arm
set -x
O=gpurun_out
mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q -x > $O/final_tests.log 2>&1; echo "rc=$?" >> $O/final_tests.log
tail -4 $O/final_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/final_smoke.log 2>&1; echo "rc=$?" >> $O/final_smoke.log
tail -3 $O/final_smoke.log
timeout 1500 python bench.py > $O/final_bench.json 2> $O/final_bench.err; echo "bench rc=$?"
tail -c 1500 $O/final_bench.json
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > $O/final_bench_reference.json 2> $O/final_bench_reference.err; echo "ref rc=$?"
cut -c1-600 $O/final_bench_reference.json
