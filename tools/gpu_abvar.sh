# development aid: stage times of the bench workload with the built library and with a variant (tools/build_variant.sh <name>), interleaved
#   gpurun -- 'bash tools/gpu_abvar.sh qpf'
run() { echo "== $1 | $2"; env $1 python tools/prof_step.py $2 2>&1 | tail -n 1 | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read().split('} ')[0]+'}')
print({k:d[k] for k in ('ms_cull','ms_field','ms_tri','ms_vertex','ms_color','ms_total')})"; }
V=PSGPU_LIB=$PWD/parsip_b200/libpsgpu_$1.so
run "A=1" ""; run "$V" ""; run "A=1" ""; run "$V" ""; run "A=1" "--slab=3/8"; run "$V" "--slab=3/8"
