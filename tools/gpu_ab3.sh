# development aid: stage times of the bench workloads + GPU test-suite
run() { echo "== $1 | $2"; env $1 python tools/prof_step.py $2 2>&1 | tail -n 1 | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read().split('} ')[0]+'}')
print({k:d[k] for k in ('mpus_evaluated','pruned_quads','ms_cull','ms_field','ms_tri','ms_vertex','ms_color','ms_total')})"; }
run "A=1" ""
run "A=1" ""
run "A=1" "--native"
run "A=1" "--slab=3/8"
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
