# development aid: cull A/B on the bench workloads (stage times printed by tools/prof_step.py)
mkdir -p gpurun_out
run() { echo "== $1 | $2"; env $1 python tools/prof_step.py $2 2>&1 | tail -n 1 | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read().split('} ')[0]+'}')
print({k:d[k] for k in ('ms_cull','ms_field','ms_tri','ms_vertex','ms_color','ms_total','pruned_quads','gpu_launches')})"; }
for w in "" "--slab=3/8" "--csgwarp"; do
  run "A=1" "$w"
  run "PSGPU_CULL_LAST_TWO_PASS=1" "$w"
  run "PSGPU_LIB=$PWD/parsip_b200/libpsgpu_noz.so" "$w"
  run "PSGPU_LIB=$PWD/parsip_b200/libpsgpu_noz.so PSGPU_CULL_LAST_TWO_PASS=1" "$w"
done > gpurun_out/cull_ab.log 2>&1
cat gpurun_out/cull_ab.log
python -m pytest tests/test_gpu_mesh.py -m gpu -q -x 2>&1 | tail -3
