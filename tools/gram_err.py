import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, oracle, oryx_b200 as ob
from als_harness import random_model
for n,f in [(70000,100),(200000,64)]:
    y = random_model(n, f, seed=n+f)*0.3
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n,dtype=np.int32), y)
        g,_ = m.gram(ob.Y)
    exp = oracle.transpose_times_self(y)
    err = np.abs(g-exp)
    print(n,f,'max abs err',err.max(),'rel to max', err.max()/np.abs(exp).max(), 'diag rel', (np.abs(np.diag(g)-np.diag(exp))/np.diag(exp)).max(), 'offdiag abs', (err - np.diag(np.diag(err))).max())
