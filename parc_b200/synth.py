"""Deterministic synthetic INPUTS for the tracking hot path (BASELINE.json configs 2, 3, 5).

There is no network for motion datasets, so the benchmark and the full-size tests run on a
seeded synthetic library of the repo humanoid: clips of 60..300 frames at
30 fps with a random-walk root, smoothed random joint angles inside half the joint limits, 0/1
contact labels and one blocky 32x32 heightfield per clip. Everything here is plain CPU torch input
generation -- no kernel, no oracle -- so the CUDA path and the CPU oracle can be fed the very same
tensors. Quaternions are (x, y, z, w).
"""
import json
import math
import os

import numpy as np
import torch

ASSET_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets")
HUMANOID_JSON = os.path.join(ASSET_DIR, "humanoid.json")


def humanoid_spec():
    """Skeleton spec of data/assets/humanoid.xml (written by KinCharModel.to_spec)."""
    with open(HUMANOID_JSON, "r") as f:
        return json.load(f)


def _exp_to_quat(e):
    ang = e.norm(dim=-1, keepdim=True)
    half = 0.5 * ang
    k = torch.where(ang > 1e-8, torch.sin(half) / ang.clamp(min=1e-8), 0.5 * torch.ones_like(ang))
    q = torch.cat([e * k, torch.cos(half)], dim=-1)
    return q / q.norm(dim=-1, keepdim=True)


def _qmul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz], dim=-1)


def _smooth(x, k):
    """moving average over time (dim 0) with window 2k+1, edge-replicated."""
    if k <= 0:
        return x
    T = x.shape[0]
    flat = x.reshape(T, -1).t().unsqueeze(1)
    pad = torch.nn.functional.pad(flat, (k, k), mode="replicate")
    out = torch.nn.functional.avg_pool1d(pad, 2 * k + 1, stride=1)
    return out.squeeze(1).t().reshape(x.shape)


def make_terrain_hf(gen, dim=32, max_boxes=6):
    """Blocky heightfield: heights 0.2 * U{-5..5} in random axis-aligned boxes over a flat floor."""
    hf = torch.zeros(dim, dim, dtype=torch.float32)
    nb = int(torch.randint(1, max_boxes + 1, (1,), generator=gen))
    for _ in range(nb):
        x0, y0 = (int(v) for v in torch.randint(0, dim - 2, (2,), generator=gen))
        w, h = (int(v) for v in torch.randint(2, dim // 2, (2,), generator=gen))
        hf[x0:x0 + w, y0:y0 + h] = 0.2 * float(torch.randint(-5, 6, (1,), generator=gen))
    return hf


def make_clips(spec, num_motions, seed=1234, min_frames=60, max_frames=300, fps=30, wrap_fraction=0.0,
               terrain_dim=32, dx=0.4):
    """-> list of clip dicts: root_pos [nf,3], root_rot [nf,4], joint_rot [nf,J-1,4], contacts [nf,J],
    fps, loop_mode (0 CLAMP / 1 WRAP), hf [dim,dim], min_point [2], dx."""
    gen = torch.Generator().manual_seed(seed)
    joints = spec["joints"]
    J = len(joints)
    clips = []
    for m in range(num_motions):
        nf = int(torch.randint(min_frames, max_frames + 1, (1,), generator=gen))
        drift = 0.02 * torch.randn(2, generator=gen)
        xy = torch.cumsum(0.03 * torch.randn(nf, 2, generator=gen) + drift, dim=0)
        z = 0.9 + 0.05 * torch.randn(nf, 1, generator=gen)
        root_pos = torch.cat([xy, _smooth(z, 3)], dim=-1)
        heading = torch.cumsum(0.05 * torch.randn(nf, generator=gen), dim=0) + 2 * math.pi * torch.rand(1, generator=gen)
        hq = torch.stack([torch.zeros(nf), torch.zeros(nf), torch.sin(0.5 * heading), torch.cos(0.5 * heading)], dim=-1)
        tilt = _exp_to_quat(_smooth(0.1 * torch.randn(nf, 3, generator=gen), 4))
        root_rot = _qmul(hq, tilt)
        root_rot = root_rot / root_rot.norm(dim=-1, keepdim=True)
        jq = torch.zeros(nf, J - 1, 4)
        jq[..., 3] = 1.0
        for j in range(1, J):
            jt = joints[j]
            if jt["type"] == "SPHERICAL":
                lim = torch.tensor(jt["limits"], dtype=torch.float32)            # [3,2]
                u = torch.rand(nf, 3, generator=gen)
                e = 0.5 * (lim[:, 0] + u * (lim[:, 1] - lim[:, 0]))
                jq[:, j - 1] = _exp_to_quat(_smooth(e, 5))
            elif jt["type"] == "HINGE":
                lim = torch.tensor(jt["limits"], dtype=torch.float32)            # [2]
                u = torch.rand(nf, 1, generator=gen)
                a = _smooth(0.5 * (lim[0] + u * (lim[1] - lim[0])), 5)
                ax = torch.tensor(jt["axis"], dtype=torch.float32)
                ax = ax / ax.norm()
                jq[:, j - 1] = torch.cat([ax * torch.sin(0.5 * a), torch.cos(0.5 * a)], dim=-1)
        con = (_smooth((torch.rand(nf, J, generator=gen) < 0.3).float(), 2) > 0.5).float()
        loop = 1 if float(torch.rand(1, generator=gen)) < wrap_fraction else 0
        hf = make_terrain_hf(gen, terrain_dim)
        centre = 0.5 * (xy.min(dim=0).values + xy.max(dim=0).values)
        # snap the terrain origin to the cell lattice like the bundled clips (min_point multiples of dx)
        min_point = torch.round((centre - 0.5 * dx * (terrain_dim - 1)) / dx) * dx
        clips.append(dict(root_pos=root_pos.contiguous(), root_rot=root_rot.contiguous(), joint_rot=jq.contiguous(),
                          contacts=con.contiguous(), fps=fps, loop_mode=loop, hf=hf, min_point=min_point, dx=dx))
    return clips


def num_padding_cells(padding, dx):
    """dm_env.py:159-166: `int(padding / dx)` after asserting that the quotient is (nearly) whole -- int() TRUNCATES, so
    padding 1.2 at dx 0.4 (2.9999999999999996) gives 2 cells; reproduced as is."""
    n = float(padding) / float(dx)
    assert (round(n) - n) < 1e-5
    return int(n)


def build_terrain_square(clips, padding_cells=1, x_offset=0.0, y_offset=0.0):
    """Blocky grid tiling all clips' terrains, same placement rule as
    DeepMimicEnv.build_terrain_square (PARC/motion_tracker/envs/ig_parkour/dm_env.py:157-316):
    ceil(sqrt(M))^2 tiles of (max_dims + 2 pad) cells, clip m's heightfield padded with its own
    minimum height, dm_motion_offsets[m, 0] = tile origin - padded clip min_point.
    -> hf [X,Y] fp32, min_point [2], dx, dm_motion_offsets [M,1,2]."""
    M = len(clips)
    dx = float(clips[0]["dx"])
    p = int(padding_cells)
    nt = int(np.ceil(np.sqrt(M)))
    dimx = max(int(c["hf"].shape[0]) for c in clips) + 2 * p
    dimy = max(int(c["hf"].shape[1]) for c in clips) + 2 * p
    first_x = -dimx * nt * dx / 2.0
    first_y = -dimy * nt * dx / 2.0
    hf = torch.zeros(dimx * nt, dimy * nt, dtype=torch.float32)
    offs = torch.zeros(M, 1, 2, dtype=torch.float32)
    m = 0
    xo = x_offset + first_x
    for i in range(nt):
        yo = y_offset + first_y
        for j in range(nt):
            if m >= M:
                break
            c = clips[m]
            padded = torch.nn.functional.pad(c["hf"], [p, p, p, p], value=float(c["hf"].min()))
            pmin = (c["min_point"] - dx * p).numpy().astype(np.float32)          # SubTerrain.pad moves min_point
            offs[m, 0, 0] = xo - float(pmin[0])
            offs[m, 0, 1] = yo - float(pmin[1])
            hf[dimx * i:dimx * i + padded.shape[0], dimy * j:dimy * j + padded.shape[1]] = padded
            m += 1
            yo += dx * dimy
        xo += dx * dimx
    min_point = torch.tensor([x_offset + first_x, y_offset + first_y], dtype=torch.float32)
    return hf, min_point, dx, offs


def xy_grid_points(dx, dy, num_x_neg, num_x_pos, num_y_neg, num_y_pos, center=(0.0, 0.0)):
    """geom_util.get_xy_grid_points (PARC/util/geom_util.py:212-224) flattened to [P, 2]: meshgrid(xs, ys, 'ij')."""
    c = torch.tensor(center, dtype=torch.float32)
    xs = torch.linspace(c[0] - dx * num_x_neg, c[0] + dx * num_x_pos, num_x_neg + num_x_pos + 1)
    ys = torch.linspace(c[1] - dy * num_y_neg, c[1] + dy * num_y_pos, num_y_neg + num_y_pos + 1)
    x, y = torch.meshgrid(xs, ys, indexing="ij")
    return torch.stack([x, y], dim=-1).reshape(-1, 2).contiguous()


def grid_ray_points(n_neg=15, n_pos=16, d=0.1):
    """geom_util.get_xy_grid_points(center=0, dx=dy=d, 15/16/15/16) -> [1024, 2] for the 32x32 config."""
    return xy_grid_points(d, d, n_neg, n_pos, n_neg, n_pos)


def cone_ray_points(dx=0.05, num_neg=2, num_pos=60, rays_neg=3, rays_pos=3, angle=0.26179938779):
    """geom_util.get_xy_points_cone with the shipped defaults (dm_env_default.yaml:144-149) -> [441,2]."""
    x = torch.linspace(-dx * num_neg, dx * num_pos, num_neg + num_pos + 1, dtype=torch.float32)
    pts = []
    for i in range(rays_neg + 1 + rays_pos):
        a = torch.ones_like(x) * (-angle * (rays_neg - i))
        c, s = torch.cos(a), torch.sin(a)
        pts.append(torch.stack([x * c - 0.0 * s, x * s + 0.0 * c], dim=-1))
    return torch.cat(pts, dim=0).contiguous()


def env_offsets(num_envs, env_spacing=5.0):
    """IGEnv env grid (ig_parkour_env.py:389-398): 2*spacing pitch, int(sqrt(N)) envs per row."""
    per_row = max(int(np.sqrt(num_envs)), 1)
    i = torch.arange(num_envs)
    out = torch.zeros(num_envs, 3, dtype=torch.float32)
    out[:, 0] = env_spacing * 2 * (i % per_row).float()
    out[:, 1] = env_spacing * 2 * (i // per_row).float()
    return out


def sim_noise(num_envs, J, D, seed=5678):
    """Seeded perturbations that turn a reference state into a 'simulated' one (section 8d)."""
    g = torch.Generator().manual_seed(seed)
    n = num_envs
    contact_on = (torch.rand(n, J, generator=g) < 0.2).float()
    forces = torch.zeros(n, J, 3)
    forces[..., 2] = contact_on * torch.abs(50.0 * torch.randn(n, J, generator=g))
    return dict(
        root_pos=0.05 * torch.randn(n, 3, generator=g),
        root_rot_exp=0.05 * torch.randn(n, 3, generator=g),
        dof_pos=0.05 * torch.randn(n, D, generator=g),
        root_vel=0.2 * torch.randn(n, 3, generator=g),
        root_ang_vel=0.2 * torch.randn(n, 3, generator=g),
        dof_vel=0.2 * torch.randn(n, D, generator=g),
        body_pos=0.01 * torch.randn(n, J, 3, generator=g),
        contact_forces=forces,
    )


def perturb_state(ref, noise, fk, dof_to_rot):
    """ref: dict root_pos/root_rot/root_vel/root_ang_vel/dof_pos/dof_vel (any device); noise from
    sim_noise (moved to that device); fk / dof_to_rot: the model's methods. -> simulated char state."""
    dev = ref["root_pos"].device
    nz = {k: v.to(dev) for k, v in noise.items()}
    root_pos = ref["root_pos"] + nz["root_pos"]
    root_rot = _qmul(ref["root_rot"], _exp_to_quat(nz["root_rot_exp"]))
    root_rot = root_rot / root_rot.norm(dim=-1, keepdim=True)
    dof_pos = ref["dof_pos"] + nz["dof_pos"]
    body_pos = fk(root_pos, root_rot, dof_to_rot(dof_pos))[0] + nz["body_pos"]
    return dict(root_pos=root_pos, root_rot=root_rot, root_vel=ref["root_vel"] + nz["root_vel"],
                root_ang_vel=ref["root_ang_vel"] + nz["root_ang_vel"], dof_pos=dof_pos,
                dof_vel=ref["dof_vel"] + nz["dof_vel"], body_pos=body_pos, contact_forces=nz["contact_forces"])


def reset_draws(num, num_motions, terrains_per_motion, lengths, weights, seed):
    """Host-side random draws of one RSI reset (what torch's generator supplies in the reference:
    multinomial, randint, rand, rand): -> motion_ids i64, terrain_ids i64, times f32, xy_noise [n,2]."""
    g = torch.Generator().manual_seed(seed)
    mids = torch.multinomial(weights.cpu(), num, replacement=True, generator=g)
    tids = torch.randint(terrains_per_motion, (num,), generator=g, dtype=torch.int64)
    times = torch.rand(num, generator=g) * lengths.cpu()[mids]
    noise = torch.rand(num, 2, generator=g) * 2.0 - 1.0
    return mids, tids, times, noise


def standin_noise(num_envs, J, D, seed=5678, fail_fraction=0.02):
    """[N, 13 + 2 D + 3 J] perturbation rows for the simulator stand-in (parc_b200_sim_standin and the same
    arithmetic in torch for the CPU arm): root pos | unit noise quaternion | root vel | root ang vel | dof pos |
    dof vel | body pos, with the scales of sim_noise. A `fail_fraction` of the envs is pushed 0.8 m off the
    reference root so that early termination keeps firing at a stationary rate."""
    nz = sim_noise(num_envs, J, D, seed)
    q = _exp_to_quat(nz["root_rot_exp"])
    root_pos = nz["root_pos"].clone()
    k = int(round(num_envs * fail_fraction))
    if k > 0:
        g = torch.Generator().manual_seed(seed + 17)
        idx = torch.randperm(num_envs, generator=g)[:k]
        root_pos[idx, 0] += 0.8
    return torch.cat([root_pos, q, nz["root_vel"], nz["root_ang_vel"], nz["dof_pos"], nz["dof_vel"],
                      nz["body_pos"].reshape(num_envs, 3 * J)], dim=1).contiguous()


def apply_standin_torch(env, noise, J, D):
    """The stand-in on an oracle TrackingEnv (torch CPU): same formulas as sim_standin_kernel. `env.model` supplies
    forward_kinematics / dof_to_rot."""
    n = noise
    env.char_root_pos[:] = env.ref_root_pos + n[:, 0:3]
    q = _qmul(env.ref_root_rot, n[:, 3:7])
    env.char_root_rot[:] = q / q.norm(dim=-1, keepdim=True)
    env.char_root_vel[:] = env.ref_root_vel + n[:, 7:10]
    env.char_root_ang_vel[:] = env.ref_root_ang_vel + n[:, 10:13]
    env.char_dof_pos[:] = env.ref_dof_pos + n[:, 13:13 + D]
    env.char_dof_vel[:] = env.ref_dof_vel + n[:, 13 + D:13 + 2 * D]
    body_pos = env.model.forward_kinematics(env.char_root_pos, env.char_root_rot, env.model.dof_to_rot(env.char_dof_pos))[0]
    env.char_rigid_body_pos[:] = body_pos + n[:, 13 + 2 * D:].reshape(-1, J, 3)
