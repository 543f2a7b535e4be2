import sys, copy, numpy as np, yaml
sys.path.insert(0,'.')
from rotors_mpc_controller_b200 import BatchedPositionNMPC
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs
params = yaml.safe_load(open(default_params_path()))
B=8192
rng=np.random.default_rng(0); x0=random_states(rng,B)
res={}
for prec,warm in (('f64',True),('f32',True),('f32',False)):
    p=copy.deepcopy(params); p['solver'].update(max_sqp_iter=30, sqp_tol=1e-4, precision=prec, ipm_warm_start=warm)
    ctl=BatchedPositionNMPC(p,batch=B)
    if not res: yref,yref_e=setpoint_refs(np.random.default_rng(1),x0,20,ctl.hover_thrust)
    import time; t=time.perf_counter(); u0,st=ctl.solve(x0,yref,yref_e); dt=time.perf_counter()-t
    s=ctl.stats(); res[(prec,warm)]=(u0.copy(),st.copy())
    r=res[('f64',True)]
    ok=(st==0)&(r[1]==0)
    rel=np.abs(u0-r[0])[ok].max()/np.abs(r[0]).max()
    print(prec,warm,'status hist',dict(zip(*np.unique(st,return_counts=True))),'facts',s['qp_iters'].mean(),'sqp',s['sqp_iters'].mean(),'max rel err vs f64',rel, f'{dt*1e3:.1f} ms')
    ctl.close()
