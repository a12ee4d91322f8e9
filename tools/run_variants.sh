#!/bin/bash
# runs the resident-only bench once per library variant in gpurun_variants/ (and the in-tree build as "base")
mkdir -p gpurun_out
for lib in base gpurun_variants/libdfu3d_*.so; do
  if [ "$lib" = base ]; then unset DFU3D_LIB; n=base; else export DFU3D_LIB=$PWD/$lib; n=$(basename $lib .so); fi
  python bench.py --steps 5 --warmup 3 --no-filters --no-e2e --no-cpu-baseline > gpurun_out/var_$n.json 2> gpurun_out/var_$n.err || tail -3 gpurun_out/var_$n.err
done
python - <<'P'
import glob, json
for f in sorted(glob.glob('gpurun_out/var_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d['kernels']
        print('%-28s %8.1f fps %7.3f ms | A %.3f B %.3f C %.3f' % (f[15:-5], d['value'], d['ms_per_step'], k['densify_a']['ms_per_step'], k['densify_b']['ms_per_step'], k['densify_c']['ms_per_step']))
    except Exception as e:
        print(f, 'failed', e)
P
