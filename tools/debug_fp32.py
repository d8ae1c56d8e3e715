import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from helpers import CONFIGS, py_tables, synth_actions
from oracle import salp_oracle
from underwater_swimmer_rl_b200 import SalpVectorEnv
N, T = 256, 400
params = CONFIGS["defaults"]
acts = synth_actions(T, N, 1, seed=11)
fw, jw = py_tables(range(9000, 9000 + N), 512, 1024)
orc = salp_oracle.Oracle(N, fw, jw, **params); orc.reset()
env = SalpVectorEnv(N, device="cuda:0", dtype="float32", autoreset=False, **params)
env.set_rng_tables(fw, jw); env.reset()
F = env.food_slots
shown = 0
for t in range(T):
    e = orc.envs
    body = np.stack([np.stack([e["px"], e["py"], e["vx"], e["vy"]], 1), np.stack([e["angle"], e["omega"], e["nozzle"], e["wv"]], 1)])
    env.body.copy_(torch.from_numpy(body.astype(np.float32)))
    ctrl = env.ctrl.cpu().numpy()
    ctrl[0, :, 0] = e["phase"] | (e["timer"] << 2) | (e["exhale_dur"] << 10) | (e["shape_kind"] << 18) | (e["shape_timer"] << 20)
    ctrl[0, :, 1] = e["steps_since_food"]; ctrl[0, :, 2] = e["food_collected"]
    mask = (e["alive"][:, :F].astype(np.int64) << np.arange(F)).sum(1)
    ctrl[0, :, 3] = np.where(mask >= 2**31, mask - 2**32, mask)
    ctrl[1, :, 0] = e["cur_food"]; ctrl[1, :, 1] = e["cur_jit"]; ctrl[1, :, 2] = (ctrl[1, :, 2] & 0x3FFFFFF) | (e["n_food"].astype(np.int64) << 26).astype(np.int32)
    env.ctrl.copy_(torch.from_numpy(ctrl)); env.refresh_rng()
    env.food.copy_(torch.from_numpy(e["food"][:, :F].transpose(1, 0, 2).astype(np.float32)))
    obs_ref, rew_ref, flags_ref = orc.step(acts[t].astype(np.float64))
    obs, rew, term, trunc, info = env.step(torch.from_numpy(acts[t]).cuda())
    o = obs.cpu().numpy()
    same = info["events"].cpu().numpy() == flags_ref
    err = np.abs(o - obs_ref) * same[:, None]
    if err.max() > 1e-3 and shown < 6:
        i, c = np.unravel_index(err.argmax(), err.shape)
        print("step", t, "env", i, "col", c, "err", err.max(), "flags", flags_ref[i], "mask", hex(int(mask[i])))
        print(" gpu", np.round(o[i, 10:], 4)); print(" ref", np.round(obs_ref[i, 10:], 4)); shown += 1
    for i in np.nonzero(flags_ref & 3)[0]:
        orc.reset_one(int(i))
print("done")
