import sys, numpy as np
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import cases
from test_gpu_parity import make
from oracle import oracleapi
g=dict(np.load('tests/golden/a16.npz'))
a,x,y,w=make('a16'); a.cache(0)
xs,rc,it,hist=a.gmres(g['gm_b'])
print('gpu', it, hist[-6:])
print('ref', int(g['gm_iters']), g['gm_hist'][-6:])
sz,deg,K,gg,sr,np_,ml,kind=cases.CASES['a16']
o=oracleapi.OracleAniso(sz,deg,K,gg,sr,np_,ml); o.set_coeff(*cases.sigma_fields(kind,x,y,sz,deg)); o.cache(0)
xo,rco,ito,ho=o.gmres(g['gm_b']); print('oracle', ito, ho[-6:])
b2=g['gm_b']*(1+1e-15*np.sin(np.arange(len(x))))
xo,rco,ito,ho=o.gmres(b2); print('oracle perturbed b', ito, ho[-6:])
