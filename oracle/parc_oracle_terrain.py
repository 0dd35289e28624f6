"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch) of PARC's motion-aware terrain post-processing
(SURVEY.md section 8f rank 4): terrain_util.compute_hf_mask_inds (PARC/util/terrain_util.py:1851-1897),
compute_hf_mask_from_inds (:1899-1906) and compute_hf_extra_vals (:1918-1947). Vectorised over frames, bodies and
points instead of the reference's triple Python loop. Imported by tests only.

Parity pin: tests/golden/ref_hf_mask.npz (tests/golden/make_golden_hf_mask.py runs the unmodified reference);
tests/test_oracle_golden_terrain.py checks this file against it."""
import torch

from . import parc_oracle as O


def body_points_world(model: O.CharModel, body_points, counts, root_pos, root_rot, joint_rot):
    """[T, Np, 3] world positions of the body sample points (:1865, :1880)."""
    body_pos, body_rot = model.forward_kinematics(root_pos, root_rot, joint_rot)
    out, start = [], 0
    T = root_pos.shape[0]
    for b, cnt in enumerate(counts):
        cnt = int(cnt)
        pts = body_points[start:start + cnt]
        start += cnt
        out.append(O.quat_rotate(body_rot[:, b].unsqueeze(1).expand(-1, cnt, -1), pts.expand(T, cnt, 3)) + body_pos[:, b].unsqueeze(1))
    return torch.cat(out, dim=1)


def compute_hf_mask_inds(model, body_points, counts, terrain: O.Terrain, root_pos, root_rot, joint_rot):
    """-> (list over frames of the unique [k,2] grid indices in lexicographic order, min_body_heights [X,Y])"""
    world = body_points_world(model, body_points, counts, root_pos, root_rot, joint_rot)
    idx = terrain.grid_index(world[..., 0:2])                                         # :1882
    X, Y = terrain.hf.shape
    key = idx[..., 0] * Y + idx[..., 1]
    mins = torch.full((X * Y,), 99999.9999)                                          # :1873-1874
    mins = mins.scatter_reduce(0, key.flatten(), world[..., 2].flatten(), reduce="amin", include_self=True)   # :1893-1896
    inds = []
    for t in range(world.shape[0]):
        k = torch.unique(key[t])                                                     # :1899 (unique rows, sorted)
        inds.append(torch.stack([k // Y, k % Y], dim=-1))
    return inds, mins.reshape(X, Y)


def compute_hf_mask_from_inds(terrain: O.Terrain, inds):
    mask = torch.zeros_like(terrain.hf, dtype=torch.bool)
    for k in inds:
        mask[k[:, 0], k[:, 1]] = True
    return mask


def compute_hf_extra_vals(model, body_points, counts, terrain: O.Terrain, root_pos, root_rot, joint_rot, z_buf=3.0, jump_buf=0.8):
    """-> (inds, hf_maxmin [X,Y,2]) (:1918-1947)"""
    inds, mins = compute_hf_mask_inds(model, body_points, counts, terrain, root_pos, root_rot, joint_rot)
    mask = compute_hf_mask_from_inds(terrain, inds)
    max_h = torch.max(root_pos[:, 2]).item()
    min_h = torch.min(terrain.hf).item()
    mm = torch.zeros(*terrain.hf.shape, 2)
    mm[..., 0] = max_h + z_buf
    mm[..., 1] = min_h - z_buf
    mm[..., 0][mask] = terrain.hf[mask]
    mm[..., 1][mask] = terrain.hf[mask]
    jump = torch.logical_and((mins - terrain.hf) >= jump_buf, mask)
    mm[..., 0][jump] = mins[jump] - jump_buf
    mm[..., 1][jump] = min_h - z_buf
    return inds, mm
