# usage: benchlib.sh <libfile> n k batch   (temporarily swaps the product library; for kernel-variant A/B runs on the GPU box)
cp mini_opt_b200/libminiopt_b200.so /tmp/keep.so
[ "$1" != "mini_opt_b200/libminiopt_b200.so" ] && cp $1 mini_opt_b200/libminiopt_b200.so
python bench.py --n $2 --k $3 --batch $4 --steps 3 --no-cpu --no-e2e 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['config']['n'], round(d['ms_per_step'],2), 'ms', round(d['roofline']['frac'],4))"
cp /tmp/keep.so mini_opt_b200/libminiopt_b200.so
