# usage: benchlib_ik.sh <libfile> links batch   (temporarily swaps the product library; IK kernel-variant A/B runs)
cp mini_opt_b200/libminiopt_b200.so /tmp/keep.so
[ "$1" != "mini_opt_b200/libminiopt_b200.so" ] && cp $1 mini_opt_b200/libminiopt_b200.so
python bench.py --workload ik --links $2 --batch $3 --steps 2 --warmup 1 --no-cpu --no-e2e 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', '$2 links', round(d['ms_per_step'],1), 'ms', round(d['nls_problems_per_sec']), 'problems/s')"
cp /tmp/keep.so mini_opt_b200/libminiopt_b200.so
