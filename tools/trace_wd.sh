# usage: trace_wd.sh <tracelib> n k batch : run the MO_WD_TRACE build and print the per-phase cycle totals of CTA 0
cp mini_opt_b200/libminiopt_b200.so /tmp/keep.so; cp $1 mini_opt_b200/libminiopt_b200.so
python bench.py --n $2 --k $3 --batch $4 --steps 1 --warmup 3 --no-cpu --no-e2e 2>&1 | grep "wd trace" | tail -2
cp /tmp/keep.so mini_opt_b200/libminiopt_b200.so
