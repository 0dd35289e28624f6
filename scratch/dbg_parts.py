import sys, torch
sys.path.insert(0, '.')
from tests import helpers as H
from parc_b200 import _lib
part = int(sys.argv[1]); ray = sys.argv[2]; n = int(sys.argv[3]); m = int(sys.argv[4]); wrap = float(sys.argv[5]); mf = int(sys.argv[6])
sc = H.Scenario(num_envs=n, num_motions=m, ray=ray, seed=77, wrap_fraction=wrap, max_frames=mf)
oenv = sc.oracle_env(); H.randomize_oracle_state(oenv, seed=31)
penv = sc.product_env("cuda:0"); H.copy_oracle_state_to_product(oenv, penv)
for it in range(2):
    penv.post_physics_step(parts=part, update_fail_rates=False)
    torch.cuda.synchronize()
    print("step", it, "ok")
print("parts", part, ray, n, m, wrap, mf, "ok")
