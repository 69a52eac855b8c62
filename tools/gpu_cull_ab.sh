# development aid: cull schedule A/B on rank slabs of blob10k@1024^3 (stage times printed by tools/prof_step.py)
mkdir -p gpurun_out
run() { echo "== $1 | $2"; env $1 python tools/prof_step.py $2 2>&1 | tail -n 1 | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read().split('} ')[0]+'}')
print({k:d[k] for k in ('ms_cull','ms_field','ms_total','pruned_quads','gpu_launches')})"; }
for w in "--slab=0/8" "--slab=3/8"; do
  run "A=1" "$w"
  run "PSGPU_CULL_LEVELS=4,1 PSGPU_CULL_CTA_MAXN=100000000" "$w"
  run "PSGPU_CULL_LEVELS=4,1 PSGPU_CULL_CTA_MAXN=100000000 PSGPU_CULL_CTA_MINLEN=256" "$w"
  run "PSGPU_CULL_CTA_MAXN=100000000 PSGPU_CULL_CTA_MINLEN=128" "$w"
done > gpurun_out/cull_ab.log 2>&1
cat gpurun_out/cull_ab.log
