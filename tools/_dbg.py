import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np
from fastbn_b200 import capi
from conftest import load_golden
t=capi.tree_from_npz('tests/golden/alarm_tree_ref.npz')
p=capi.Plan(t,device=0)
g=load_golden('alarm_edge_ref')
lab,post,_=p.run(g['ev_off'],g['ev_vars'],g['ev_vals'])
d=np.abs(post-g['post'])
print('nan ours',np.isnan(post).sum(axis=1),'nan ref',np.isnan(g['post']).sum(axis=1))
print('maxabs per case',np.nanmax(d,axis=1))
print(lab,g['labels'])
i=int(np.nanargmax(np.nanmax(d,axis=1)))
print(i, post[i][:12], g['post'][i][:12])
