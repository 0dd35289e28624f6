#!/usr/bin/env python
"""BASELINE config 4: kinematic motion optimizer batch, 1,024 motions x 300 frames, loss forward + backward
(FK + terrain penetration/contact SDF + temporal regularisers) through parc_b200.motion_optimization.

    python bench_optimizer.py [--motions 1024] [--frames 300] [--steps 10] [--warmup 3] [--cpu-motions 2]

Prints one JSON line: motions/s and ms per value-and-gradient evaluation on the GPU (CUDA events), the
per-kernel split, and the CPU oracle port (torch autograd) on a bounded sample of the same batch. This is
a secondary measurement; the driver's bench.py stays on config 2."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LAST_EVALS = 0
KIN_GEN = dict(w_root_pos=1.0, w_root_rot=10.0, w_joint_rot=1.0, w_smoothness=10.0, w_penetration=1000.0, w_contact=1000.0,
               w_sliding=10.0, w_body_constraints=1000.0, w_jerk=1000.0, max_jerk=1000.0)
NAMES = ("src_root_pos", "src_root_rot_quat", "src_joint_rot", "src_body_vels", "src_body_rot_vels", "contacts")


def build_batch(model, B, T, seed, dev, distinct=True):
    """Synthetic clips (SURVEY 8d generator) with their own 32x32 terrains; source features by the CUDA FK."""
    from parc_b200 import synth
    gen = B if distinct else min(B, 64)
    clips = synth.make_clips(synth.humanoid_spec(), gen, seed=seed, min_frames=T, max_frames=T, wrap_fraction=0.0)
    rp = torch.stack([c["root_pos"] for c in clips]).to(dev)
    rr = torch.stack([c["root_rot"] for c in clips]).to(dev)
    jr = torch.stack([c["joint_rot"] for c in clips]).to(dev)
    hf0 = torch.stack([torch.as_tensor(c["hf"], dtype=torch.float32) for c in clips]).to(dev)
    mp0 = torch.stack([torch.as_tensor(c["min_point"], dtype=torch.float32) for c in clips]).to(dev)
    # ground the character: root height = terrain height under the root + 0.9 m (the generator's clips float at a
    # fixed height over blocky terrain, which would bury half the body in the taller blocks)
    dx0 = float(clips[0]["dx"])
    ij = torch.round((rp[..., :2] - mp0[:, None, :]) / dx0).long().clamp(0, hf0.shape[1] - 1)
    ground = hf0[torch.arange(gen, device=dev)[:, None], ij[..., 0], ij[..., 1]]
    rp = rp.clone()
    rp[..., 2] = ground + 0.9
    bp, br = model.forward_kinematics(rp, rr, jr)
    d = dict(src_root_pos=rp, src_root_rot_quat=rr, src_joint_rot=jr, src_body_vels=bp[:, 1:] - bp[:, :-1])
    dq = br[:, 1:], br[:, :-1]
    # quat_diff_angle on the host side of the harness (input preparation, not the timed path)
    q1, q0 = dq[1], dq[0]
    conj = torch.cat([-q0[..., :3], q0[..., 3:]], dim=-1)
    w = q1[..., 3] * conj[..., 3] - (q1[..., :3] * conj[..., :3]).sum(-1)
    v = q1[..., 3:] * conj[..., :3] + conj[..., 3:] * q1[..., :3] + torch.cross(q1[..., :3], conj[..., :3], dim=-1)
    ln = v.norm(dim=-1)
    d["src_body_rot_vels"] = torch.where(ln > 1e-5, 2.0 * torch.atan2(ln, w.abs()), torch.zeros_like(ln))
    d["contacts"] = torch.stack([c["contacts"] for c in clips]).to(dev)
    d["hf"] = torch.stack([torch.as_tensor(c["hf"], dtype=torch.float32) for c in clips]).to(dev)
    d["min_point"] = torch.stack([torch.as_tensor(c["min_point"], dtype=torch.float32) for c in clips]).to(dev)
    g = torch.Generator().manual_seed(seed + 1)
    d["tgt_root_pos"] = rp + (0.02 * torch.randn(gen, T, 3, generator=g) + torch.tensor([0.0, 0.0, -0.02])).to(dev)
    d["tgt_root_rot"] = model_rot_to_exp(rr) + 0.03 * torch.randn(gen, T, 3, generator=g).to(dev)
    d["tgt_joint_dof"] = model.rot_to_dof(jr) + 0.05 * torch.randn(gen, T, 28, generator=g).to(dev)
    rep = (B + gen - 1) // gen
    d = {k: v.repeat(rep, *([1] * (v.dim() - 1)))[:B].contiguous() for k, v in d.items()}
    return d, float(clips[0]["dx"])


def model_rot_to_exp(q):
    s = torch.where(q[..., 3:] < 0, -1.0, 1.0)
    q = q * s
    ln = q[..., :3].norm(dim=-1, keepdim=True)
    ang = 2.0 * torch.atan2(ln, q[..., 3:])
    axis = q[..., :3] / ln.clamp(min=1e-6)
    return torch.where(ln > 1e-5, axis * ang, torch.zeros_like(axis))


def run(B=1024, T=300, steps=10, warmup=3, cpu_motions=2, dev=None, peak_gbs=None):
    """One measurement of BASELINE configs[3]; returns the JSON-able dict (bench.py's `extra.config4` calls this too)."""
    assert torch.cuda.is_available(), "needs a CUDA device (no CPU fallback on the product path)"
    dev = dev or torch.device("cuda", 0)
    from parc_b200 import KinCharModel, geom_util, synth
    from parc_b200.motion_optimization import motion_loss_batched
    model = KinCharModel(dev)
    model.load_char_file(synth.HUMANOID_JSON)
    pts = geom_util.get_char_point_samples(model)
    d, dx = build_batch(model, B, T, 77, dev, distinct=True)
    dxdy = torch.tensor([dx, dx])

    def step():
        p = [d[k].clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
        loss, terms = motion_loss_batched(p[0], p[1], p[2], *[d[k] for k in NAMES], d["hf"], d["min_point"], dxdy, pts, model, **KIN_GEN)
        loss.sum().backward()
        return loss, terms, p

    global LAST_EVALS
    LAST_EVALS = max(warmup, 3) + steps         # evaluations this call makes (profiles/opt_cells.py divides its counters by it)
    for _ in range(max(warmup, 3)):
        step()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss, terms, p = step()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    # CPU oracle port on a bounded sample (test infrastructure used as the reported baseline only)
    cpu = None
    if cpu_motions > 0:
        from oracle import parc_oracle_opt as OO
        from tests import helpers as H
        om = H.oracle_model()
        bp = torch.cat([x.cpu() for x in pts])
        counts = [int(x.shape[0]) for x in pts]
        w = [KIN_GEN[k] for k in ("w_root_pos", "w_root_rot", "w_joint_rot", "w_smoothness", "w_penetration", "w_contact", "w_sliding",
                                  "w_body_constraints", "w_jerk", "max_jerk")]
        torch.set_num_threads(os.cpu_count() or 1)
        t0 = time.perf_counter()
        for b in range(cpu_motions):
            po = [d[k][b].cpu().clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
            lo, _ = OO.motion_loss(om, bp, counts, d["hf"][b].cpu(), d["min_point"][b].cpu(), dxdy, po[0], po[1], po[2],
                                   *[d[k][b].cpu() for k in NAMES], w)
            lo.backward()
        dt = (time.perf_counter() - t0) / cpu_motions
        cpu = {"value": 1.0 / dt, "unit": "motions/s", "cores": os.cpu_count(), "kind": "port",
               "sample": "%d of the %d motions, torch CPU autograd, %.2f s per motion fwd+bwd" % (cpu_motions, B, dt)}
    frames = B * T
    alg_bytes = frames * 266 * 4            # params + grads + source pose + contacts + source body vels per frame
    if peak_gbs is None:
        peak_gbs = 6550.7
        mp = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(mp):
            peak_gbs = float(json.load(open(mp))["hbm_gbs"])
    achieved = alg_bytes / (ms * 1e-3) / 1e9
    # fp32 side: cells evaluated per frame were counted once with a -DPARC_OPT_COUNT_CELLS build (profiles/opt_cells.py);
    # one evaluated cell = both column distances (solid and air) = 2 x ~20 FLOP
    fp32 = None
    cells_json = os.path.join(ROOT, "profiles", "opt_sdf_cells.json")
    if os.path.exists(cells_json) and (B, T) == (1024, 300):
        cj = json.load(open(cells_json))
        flop = frames * cj["cells_per_frame"] * 40.0
        peak_tf = 148 * 128 * 2 * 1.965e9 / 1e12                       # FP32 FMA lanes x 2 x SM clock
        fp32 = {"cells_per_scan": cj["cells_per_scan"], "scans_per_frame": cj["scans_per_frame"], "cells_per_frame": cj["cells_per_frame"],
                "flop_per_eval": flop, "achieved": flop / (ms * 1e-3) / 1e12, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": flop / (ms * 1e-3) / 1e12 / peak_tf,
                "survey_formula_flop_per_eval": frames * 304 * cj["cells_per_scan"] * 2 * 20 * 3,
                "source": "profiles/opt_sdf_cells.json (counting build, one evaluation of this workload)"}
    return {"metric": "optimizer loss fwd+bwd, motions/s (1,024 x 300 frames)", "value": B / (ms * 1e-3), "unit": "motions/s",
            "ms_per_eval": ms, "frames_per_s": frames / (ms * 1e-3),
            "config": {"workload": "BASELINE configs[3]: %d distinct synthetic motions x %d frames, own 32x32 terrains" % (B, T), "motions": B,
                       "frames": T, "body_points": int(sum(x.shape[0] for x in pts)), "terrain": list(d["hf"].shape[1:])},
            "roofline": {"bound": "fp32 ALU / shared memory (SURVEY 8d); HBM floor shown", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s",
                         "frac": achieved / peak_gbs, "algorithmic_bytes_per_eval": alg_bytes, "hbm_floor_ms": alg_bytes / (peak_gbs * 1e9) * 1e3,
                         "fp32": fp32},
            "cpu_baseline": cpu,
            "stats": {"mean_loss": float(loss.mean()), "mean_penetration": float(terms[:, 4].mean())}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--motions", type=int, default=1024)
    ap.add_argument("--frames", type=int, default=300)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu-motions", type=int, default=2)
    args = ap.parse_args()
    print(json.dumps(run(args.motions, args.frames, args.steps, args.warmup, args.cpu_motions)))


if __name__ == "__main__":
    main()
