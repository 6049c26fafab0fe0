# A/B on the GPU box: parity tests of the DMMA shapes, then timings of the warp kernel with 1 and 2 warps per problem
# (CTA=1 adds the CTA-per-problem kernel)
for w in 1 2; do MO_QP_WD_WARPS=$w timeout 300 python -m pytest tests/test_gpu_qp.py -m gpu -x -q -k "box_synth_vs_oracle or dmma_shapes or box_synth_vs_reference" 2>&1 | tail -2; done
for cfg in "64 16 65536" "32 8 262144" "16 8 524288"; do set -- $cfg
  for w in 1 2; do
    MO_QP_WD_WARPS=$w timeout 200 python bench.py --n $1 --k $2 --batch $3 --steps 3 --no-cpu --no-e2e 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n=$1 warps=$w', round(d['ms_per_step'],2), 'ms frac', round(d['roofline']['frac'],4), 'conv', round(d['converged_fraction'],5), round(d['config']['newton_steps_per_qp'],4))"
  done
  [ -n "$CTA" ] && MO_QP_DM_KERNEL=cta timeout 200 python bench.py --n $1 --k $2 --batch $3 --steps 3 --no-cpu --no-e2e 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n=$1 cta', round(d['ms_per_step'],2), 'ms frac', round(d['roofline']['frac'],4))"
done
