#!/usr/bin/env python
"""MDM dataset sampler, sample_motion_data on a batch of windows.

    python bench_sampler.py [--samples 65536] [--motions 256] [--steps 20] [--warmup 3] [--cpu-samples 2048]

Workload: synthetic library (SURVEY 8d generator, each clip with its own 32x32 terrain), the shipped sampler
config (0.5 s windows at 30 fps = 15 frames, 31x31 local heightmap at 0.2 m, target info on). Prints one JSON line:
windows/s on the GPU (CUDA events, inputs resident), the roofline of the two launches against MEASURED_PEAKS.json,
and the CPU oracle port (oracle/parc_oracle_sampler.py, all host threads) on a bounded sample. The reference's
training batch is 64 windows per iteration (parc_1_train_gen_default.yaml:6); the kernels are sized for the
offline / large-batch use (dataset statistics, get_motion_sequences_for_id) where one call covers 1e4-1e5 windows.
Secondary measurement: the driver's bench.py stays on BASELINE config 2."""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def sampler_cfg(lib, dev):
    from parc_b200 import synth
    return {"device": dev, "char_file": synth.HUMANOID_JSON, "motion_lib_file": lib,
            "features": {"canonicalize_samples": True, "frame_components": ["ROOT_POS", "ROOT_ROT", "JOINT_POS", "JOINT_ROT", "CONTACTS"]},
            "sequence_fps": 30, "sequence_duration": 0.5, "autoregressive": True, "num_prev_states": 2, "future_pos_noise_scale": 0.05,
            "relative_z_style": "RELATIVE_TO_ROOT", "hf_augmentation_mode": "NONE",
            "heightmap": {"horizontal_scale": 0.2, "local_grid": {"num_x_neg": 10, "num_x_pos": 20, "num_y_neg": 15, "num_y_pos": 15},
                          "max_h": 3.0},
            "use_hf_augmentation": False, "angle_noise_scale": 0.01, "pos_noise_scale": 0.01, "future_window_min": 0.4,
            "future_window_max": 1.5}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=65536)
    ap.add_argument("--motions", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu-samples", type=int, default=2048)
    ap.add_argument("--augment", action="store_true", help="also time the call with the heightfield augmenter (MAXPOOL_AND_BOXES, "
                    "parc_1_train_gen_default.yaml values), records drawn once beforehand")
    args = ap.parse_args()
    assert torch.cuda.is_available(), "needs a CUDA device (no CPU fallback on the product path)"
    dev = torch.device("cuda", 0)
    from parc_b200 import KinCharModel, MotionLib, SubTerrain, synth
    from parc_b200.motion_sampler import MDMHeightfieldContactMotionSampler
    model = KinCharModel(dev)
    model.load_char_file(synth.HUMANOID_JSON)
    clips = synth.make_clips(synth.humanoid_spec(), args.motions, seed=99, wrap_fraction=0.05)
    pc = []
    for c in clips:
        hf = torch.as_tensor(c["hf"], dtype=torch.float32)
        t = SubTerrain(x_dim=hf.shape[0], y_dim=hf.shape[1], dx=float(c["dx"]), dy=float(c["dx"]), min_x=float(c["min_point"][0]),
                       min_y=float(c["min_point"][1]), device=dev)
        t.hf = hf.to(dev).contiguous()
        pc.append(dict(root_pos=c["root_pos"], root_rot=c["root_rot"], joint_rot=c["joint_rot"], contacts=c["contacts"], fps=c["fps"],
                       loop_mode=c["loop_mode"], terrain=t))
    lib = MotionLib.from_clips(pc, model, dev, contact_info=True)
    s = MDMHeightfieldContactMotionSampler(sampler_cfg(lib, dev))
    n, T, J = args.samples, s.get_seq_len(), 15
    g = torch.Generator().manual_seed(5)
    ids = torch.randint(0, args.motions, (n,), generator=g).to(dev)
    starts = (torch.rand(n, generator=g).to(dev) * (lib.get_motion_length(ids) - 0.5)).clamp(min=0.0)
    ft = starts + 0.4 + torch.rand(n, generator=g).to(dev)
    z = torch.randn(n, 3, generator=g).to(dev)

    def step():
        return s.sample_motion_data(motion_ids=ids, motion_start_times=starts, future_times=ft, future_pos_noise=z)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    # algorithmic bytes per window: outputs T*(3+4+3(J-1)+4(J-1)+C) + gx*gy + 7 floats; inputs 2 adjacent 320 B records per frame
    out_f = T * (3 + 4 + 7 * (J - 1) + J) + 31 * 31 + 7 + 12
    b_win = 4 * out_f + T * 640 + 20
    peak = 6650.0
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = float(json.load(open(pk))["hbm_gbs"])
    achieved = n * b_win / (ms * 1e-3) / 1e9
    aug = None
    if args.augment:
        # what the augmenter reads per clip: hf_maxmin (here hf +- 0.25 m) and the per-frame cells under the body (the product's own
        # compute_hf_mask_inds, the function scripts/parc_2_kin_gen.py runs when it writes a clip)
        from parc_b200 import geom_util
        from parc_b200.file_io import MotionFrames
        from parc_b200.terrain import compute_hf_mask_inds
        pts = geom_util.get_char_point_samples(model)
        masks = []
        for k, c in enumerate(pc):
            t_ = c["terrain"]
            t_.hf_maxmin = torch.stack([t_.hf + 0.25, t_.hf - 0.25], dim=-1).contiguous()
            fr = MotionFrames(root_pos=c["root_pos"].to(dev), root_rot=c["root_rot"].to(dev), joint_rot=c["joint_rot"].to(dev))
            masks.append([x.cpu() for x in compute_hf_mask_inds(fr, t_, model, pts)[0]])
        lib._hf_mask_inds = masks
        cfg_a = sampler_cfg(lib, dev)
        cfg_a.update(use_hf_augmentation=True, hf_augmentation_mode="MAXPOOL_AND_BOXES", max_num_boxes=4, box_min_len=2, box_max_len=12,
                     hf_maxpool_chance=0.15, hf_max_maxpool_size=10, hf_change_height_chance=0.1)
        sa = MDMHeightfieldContactMotionSampler(cfg_a)
        t0 = time.perf_counter()
        rec = sa._draw_box_aug_records(n)
        draw_s = time.perf_counter() - t0
        rec = (rec[0].to(dev), rec[1].to(dev))

        def step_a():
            return sa.sample_motion_data(motion_ids=ids, motion_start_times=starts, future_times=ft, future_pos_noise=z, hf_aug_records=rec)

        for _ in range(3):
            step_a()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.steps):
            step_a()
        e1.record()
        torch.cuda.synchronize()
        ms_a = e0.elapsed_time(e1) / args.steps
        aug = {"ms_per_call": ms_a, "augment_ms": ms_a - ms, "windows_per_s": n / (ms_a * 1e-3),
               "bytes_per_window_augment": 31 * 31 * 4 * 2 + 4 * (8 + 29), "host_draw_s": draw_s,
               "note": "MAXPOOL_AND_BOXES with parc_1_train_gen_default.yaml chances; augment_ms = the third launch (mask bitmap, bounds, pools, "
                       "boxes in shared memory); host_draw_s = Python random / torch draws of the records for all windows, outside the timed calls"}
    cpu = None
    if args.cpu_samples > 0:
        from oracle import parc_oracle as O
        from oracle import parc_oracle_sampler as OS
        from tests import helpers as H
        torch.set_num_threads(os.cpu_count() or 1)
        om = H.oracle_model()
        olib = O.Library(om, [dict(c, terrain=O.Terrain(c["hf"], c["min_point"], [c["dx"], c["dx"]])) for c in clips])
        ocfg = OS.SamplerConfig()
        k = args.cpu_samples
        a = (ids[:k].cpu(), starts[:k].cpu(), ft[:k].cpu(), z[:k].cpu())
        OS.sample_motion_data(om, olib, ocfg, a[0][:64], a[1][:64], future_t=a[2][:64], future_noise=a[3][:64])
        t0 = time.perf_counter()
        OS.sample_motion_data(om, olib, ocfg, a[0], a[1], future_t=a[2], future_noise=a[3])
        dt = time.perf_counter() - t0
        cpu = {"value": k / dt, "unit": "windows/s", "cores": os.cpu_count(), "kind": "port",
               "sample": "%d of the %d windows, torch CPU eager (per-sample terrain lookup loop as upstream), %.2f s" % (k, n, dt)}
    print(json.dumps({"metric": "MDM sampler sample_motion_data, windows/s (15 frames + 31x31 heightmap + target)", "value": n / (ms * 1e-3),
                      "unit": "windows/s", "ms_per_call": ms, "frames_per_s": n * T / (ms * 1e-3),
                      "config": {"workload": "SURVEY 8f rank 1", "windows": n, "clips": args.motions, "seq_len": T, "grid": [31, 31]},
                      "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                                   "bytes_per_window": b_win}, "augmenter": aug, "cpu_baseline": cpu}))


if __name__ == "__main__":
    main()
