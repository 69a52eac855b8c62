import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes
root=scenes.blob1k(); fb=bt.flatten(bt.binarize(root)); lo,hi=ps.prepare_bboxes(fb); cs=scenes.cellsize_for(lo,hi,512)
ctx=ps.Context(0); ctx.upload(ps.linearize(bt.flatten(root), ps.SEM_SCALAR, ps.LIN_REASSOCIATE))
m=ctx.polygonize(ps.Params.make(cs,lo,hi))
tri=m['tri'].astype(np.int64)
e=np.concatenate([tri[:,[0,1]],tri[:,[1,2]],tri[:,[2,0]]]); e.sort(axis=1)
key=e[:,0]*(m['n_vertices']+1)+e[:,1]
u,c=np.unique(key,return_counts=True)
bad=u[c==1]; v=np.unique(np.concatenate([bad//(m['n_vertices']+1), bad%(m['n_vertices']+1)]))
p=m['pos'][v]; rel=(p-lo)/cs; dims=np.array(ps.lattice_dims(cs,lo,hi))*7
print('open edges',len(bad),'verts',len(v)); print('min rel',rel.min(0),'max rel',rel.max(0),'dims',dims)
onb=((rel<1e-3)|(np.abs(rel-dims)<1e-3)).any(1); print('on domain boundary:',onb.sum(),'of',len(v))
print(rel[~onb][:10])
bad4=u[c>2]; print('pinched',len(bad4), c[c>2][:12])
