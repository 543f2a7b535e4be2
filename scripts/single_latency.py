"""configs[0] latency probe: ONE vehicle through PositionNMPC.solve (host buffers) in closed loop,
hover (0,0,1) -> set-point (2,0,1.5); prints p50 / p99 of the call and the final position.
NMPC_B200_LIB selects the library build (A/B runs)."""
import copy, os, sys, time
import numpy as np, torch, yaml
sys.path.insert(0, '.')
from rotors_mpc_controller_b200 import ClosedLoop, Plant, PositionNMPC, ReferenceGenerator
from rotors_mpc_controller_b200.params import default_params_path

params = yaml.safe_load(open(default_params_path()))
tag = os.environ.get('NMPC_B200_LIB', 'default').split('/')[-1]
p1 = copy.deepcopy(params)
one = PositionNMPC(p1, device=0)
ref = ReferenceGenerator(p1['reference'])
ref.update_hover_thrust(one.hover_thrust)
ref.set_target(position=[2.0, 0.0, 1.5], yaw=0.0)
horizon = ref.build_horizon(one.horizon, one.dt)
plant = ClosedLoop(one, Plant(dt=one.dt, substeps=5))
for rep in range(3):
    x = np.array([0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0], float)
    plant.reset(x[None])
    lat = []
    for step in range(200):
        state = {'position': x[0:3], 'velocity': x[3:6], 'quaternion': x[6:10], 'body_rates': x[10:13]}
        t0 = time.perf_counter()
        u0, status = one.solve(state, horizon)
        lat.append(time.perf_counter() - t0)
        thrust, _ = one.command()
        x = plant.plant_step(thrust[None]).cpu().numpy()[0]
    lat = np.array(lat[20:]) * 1e6
    print(f'{tag:12s} rep {rep}: p50 {np.median(lat):7.1f} us  p99 {np.percentile(lat, 99):7.1f} us  min {lat.min():7.1f}  '
          f'final {x[0]:.9f} {x[1]:.3e} {x[2]:.9f} status {int(status)}', flush=True)
one.close()
