"""bench.py -- env-steps/s of the fused obs + reward + termination + RSI-reset step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1]): 65,536 envs per GPU, 1,000-clip synthetic library of the repo
humanoid, 32x32 local heightmap observations (obs row = 1,895 floats). A "step" is one pass of
IGEnv._post_physics_step (heights, time, reference motion, observation, reward, termination),
the fail-rate EMA and the reference-state-init reset of the envs that finished. Envs shard by
env index over ranks with no data-path collective ("weak" scaling: 65,536 envs on every GPU);
the only exchange is one all-reduce of reward / termination statistics at the end.

Own arm prints ONE JSON line with value (inputs resident in HBM), e2e (host buffers, H2D/D2H in
the timed region), roofline of the dominant kernel and a cpu_baseline (the CPU oracle port, rank 0,
N=1 only). `--impl reference` times the CPU port alone on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ENVS_PER_GPU = 65536
NUM_MOTIONS = 1000
CPU_SAMPLE_ENVS = 4096
METRIC = "env-steps/s (obs+reward+RSI) at 65,536 envs/GPU; achieved HBM GB/s vs peak"
UNIT = "env-steps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", type=str, default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU, help="envs per GPU (default = the BASELINE config)")
    ap.add_argument("--motions", type=int, default=NUM_MOTIONS)
    ap.add_argument("--cpu-envs", type=int, default=CPU_SAMPLE_ENVS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fail-fraction", type=float, default=0.02,
                    help="share of the envs the simulator stand-in pushes off their reference every step (they terminate and are "
                         "reset): 0.02 is the stationary BASELINE configs[1] workload, 0.1 with --envs 262144 is configs[4] (reset-heavy)")
    ap.add_argument("--chunks", type=int, default=6, help="row ranges of the host-fed step (upload of range c+1 overlaps the step of range c)")
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# workload construction (product side: synthetic inputs + CUDA classes, no oracle)
# ------------------------------------------------------------------------------------------------
def build_world(num_motions, seed=1234):
    from parc_b200 import synth
    clips = synth.make_clips(synth.humanoid_spec(), num_motions, seed=seed, wrap_fraction=0.05)
    hf, min_point, dx, dm_offsets = synth.build_terrain_square(clips, padding_cells=1)
    return dict(clips=clips, hf=hf, min_point=min_point, dx=dx, dm_offsets=dm_offsets, ray=synth.grid_ray_points())


def build_gpu_env(world, num_envs, first_env, total_envs, device):
    """Rank-local env over global env ids [first_env, first_env + num_envs) of a total_envs grid."""
    from parc_b200 import KinCharModel, MotionLib, SubTerrain, synth
    from parc_b200.tracking_env import TrackingEnv, default_env_config
    model = KinCharModel(device)
    model.load_char_file(synth.HUMANOID_JSON)
    pc = [dict(root_pos=c["root_pos"], root_rot=c["root_rot"], joint_rot=c["joint_rot"], contacts=c["contacts"], fps=c["fps"],
               loop_mode=c["loop_mode"], terrain=None) for c in world["clips"]]
    lib = MotionLib.from_clips(pc, model, device, contact_info=True)
    hf = world["hf"]
    terrain = SubTerrain(x_dim=hf.shape[0], y_dim=hf.shape[1], dx=world["dx"], dy=world["dx"],
                         min_x=float(world["min_point"][0]), min_y=float(world["min_point"][1]), device=device)
    terrain.hf = hf.to(device).contiguous()
    offs = synth.env_offsets(total_envs)[first_env:first_env + num_envs]
    env = TrackingEnv(default_env_config(), num_envs, device, model, lib, terrain, world["dm_offsets"], world["ray"],
                      env_offsets=offs, seed=20260 + first_env)
    return env


def init_gpu_state(env, first_env, seed=4321):
    """RSI-reset every env with seeded draws, scatter them over their episodes, and install a
    perturbed 'simulated' state around the reference pose."""
    from parc_b200 import synth
    N, dev = env.get_num_envs(), env._device
    lib = env._mlib
    draws = synth.reset_draws(N, lib.num_motions(), 1, lib._motion_lengths, lib._motion_weights, seed=seed + first_env)
    env.reset(torch.arange(N, device=dev), draws=draws, refresh_obs=False)
    g = torch.Generator().manual_seed(seed + 7 + first_env)
    env._timestep_buf.copy_(torch.randint(0, 100, (N,), generator=g, dtype=torch.int32).to(dev))
    env._time_buf.copy_(env._timestep_buf.float() * np.float32(1.0 / 30.0))
    env.post_physics_step(parts=2, update_fail_rates=False)      # PARC_STEP_REF: reference state at the current time
    nz = synth.sim_noise(N, env._J, env._D, seed=5678 + first_env)
    ref = dict(root_pos=env._ref_root_pos, root_rot=env._ref_root_rot, root_vel=env._ref_root_vel,
               root_ang_vel=env._ref_root_ang_vel, dof_pos=env._ref_dof_pos, dof_vel=env._ref_dof_vel)
    sim = synth.perturb_state(ref, nz, env._kin_char_model.forward_kinematics, env._kin_char_model.dof_to_rot)
    env._char_root_pos.copy_(sim["root_pos"]); env._char_root_rot.copy_(sim["root_rot"])
    env._char_root_vel.copy_(sim["root_vel"]); env._char_root_ang_vel.copy_(sim["root_ang_vel"])
    env._char_dof_pos.copy_(sim["dof_pos"]); env._char_dof_vel.copy_(sim["dof_vel"])
    env._char_rigid_body_pos.copy_(sim["body_pos"]); env._char_contact_forces.copy_(sim["contact_forces"])
    torch.cuda.synchronize(dev)


def algorithmic_bytes(env):
    """N * B_env(P) + B_lib, B_env = 4,940 + 4 P bytes,
    B_lib = min(N * 4,504, 448 F) + 4 X Y."""
    N = env.get_num_envs()
    P = int(env._ray_xy_points.shape[0])
    F = int(env._mlib._frame_root_pos.shape[0])
    X, Y = env._terrain.hf.shape
    b_env = 4940 + 4 * P
    b_lib = min(N * 4504, 448 * F) + 4 * X * Y
    return N * b_env + b_lib, b_env


LAUNCHES_PER_STEP = 6   # device-resident loop: sim_standin, env_step, fail_rate_ema, env_reset_sampled, env_step(reset rows), done-count clear


# ------------------------------------------------------------------------------------------------
# CPU leg: the oracle port (oracle/parc_oracle.py) on the host cores -- checker code used as the
# reported baseline only, never on the product path
# ------------------------------------------------------------------------------------------------
def run_cpu_port(world, num_envs, steps, warmup, threads, fail_fraction=0.02):
    from oracle import parc_oracle as O
    from parc_b200 import synth
    from parc_b200.tracking_env import default_env_config
    torch.set_num_threads(threads)
    spec = synth.humanoid_spec()
    m = O.CharModel(spec)
    lib = O.Library(m, [dict(c, terrain=None) for c in world["clips"]])
    terrain = O.Terrain(world["hf"], world["min_point"], [world["dx"], world["dx"]])
    ec = default_env_config()
    cfg = O.EnvConfig(m, rand_root_pos_offset_scale=ec["rand_root_pos_offset_scale"])
    env = O.TrackingEnv(m, lib, terrain, world["dm_offsets"], cfg, num_envs, synth.env_offsets(num_envs), world["ray"])
    N = num_envs
    draws = synth.reset_draws(N, lib.num_motions(), 1, lib.lengths, lib.motion_weights, seed=4321)
    env.reset(torch.arange(N), *draws)
    g = torch.Generator().manual_seed(4328)
    env.timestep_buf[:] = torch.randint(0, 100, (N,), generator=g, dtype=torch.int32)
    env.time_buf[:] = env.timestep * env.timestep_buf
    env.update_ref_motion()
    ref = dict(root_pos=env.ref_root_pos, root_rot=env.ref_root_rot, root_vel=env.ref_root_vel,
               root_ang_vel=env.ref_root_ang_vel, dof_pos=env.ref_dof_pos, dof_vel=env.ref_dof_vel)
    sim = synth.perturb_state(ref, synth.sim_noise(N, m.J, m.D, seed=5678), m.forward_kinematics, m.dof_to_rot)
    env.char_root_pos[:], env.char_root_rot[:] = sim["root_pos"], sim["root_rot"]
    env.char_root_vel[:], env.char_root_ang_vel[:] = sim["root_vel"], sim["root_ang_vel"]
    env.char_dof_pos[:], env.char_dof_vel[:] = sim["dof_pos"], sim["dof_vel"]
    env.char_rigid_body_pos[:], env.char_contact_forces[:] = sim["body_pos"], sim["contact_forces"]

    noise = synth.standin_noise(N, m.J, m.D, seed=5678, fail_fraction=fail_fraction)
    resets = []

    def one_step(i):
        synth.apply_standin_torch(env, noise, m.J, m.D)      # the same simulator stand-in as the GPU arm
        env.post_physics_step()
        ids = torch.nonzero(env.done_buf != O.DONE_NULL).flatten()
        resets.append(int(ids.numel()))
        if ids.numel() > 0:
            d = synth.reset_draws(ids.numel(), lib.num_motions(), 1, lib.lengths, env.sampling_weights(), seed=1000 + i)
            env.reset(ids, *d)
            env.refresh_obs_after_reset(ids)

    for i in range(warmup):
        one_step(i)
    times = []
    for i in range(steps):
        t0 = time.perf_counter()
        one_step(warmup + i)
        times.append(time.perf_counter() - t0)
    total = float(np.sum(times))
    return {"value": N * steps / total, "median_s": float(np.median(times)), "min_s": float(np.min(times)), "total_s": total,
            "resets_per_step": float(np.mean(resets[warmup:]))}


def cpu_baseline_block(world, args, steps, warmup):
    cores = os.cpu_count() or 1
    r = run_cpu_port(world, args.cpu_envs, steps, warmup, cores, args.fail_fraction)
    return {"value": r["value"], "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d envs x %d steps (+%d warm-up) of the same %d-clip / 32x32-heightmap workload, torch CPU eager, "
                      "%d threads; median %.3f s/step, min %.3f s/step; %.1f resets/step"
                      % (args.cpu_envs, steps, warmup, args.motions, cores, r["median_s"], r["min_s"], r["resets_per_step"])}, r


def run_reference_arm(args):
    """`--impl reference`: the reference's CPU path for this workload = the oracle port (the
    reference itself is a Python tree under /root/reference that does not travel to the GPU box)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = build_world(args.motions)
    steps, warmup = max(args.steps, 1), max(args.warmup, 1)
    # bounded: ~0.25 s/step at 4,096 envs on 8 threads -> cap the sample so the run stays within minutes
    steps_cpu, warm_cpu = min(steps, 40), min(warmup, 5)
    cb, r = cpu_baseline_block(world, args, steps_cpu, warm_cpu)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps_cpu,
            "warmup": warm_cpu, "ms_per_step": 1e3 * r["total_s"] / steps_cpu, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "65,536-env / 1,000-clip / 32x32-heightmap tracking step (same stand-in + step + reset as the b200 "
                                   "arm); CPU arm runs a %d-env sample per step"
                                   % args.cpu_envs, "envs_per_step": args.cpu_envs, "clips": args.motions, "ray_points": 1024,
                       "obs_dim": 1895},
            "cpu_baseline": cb,
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch.distributed as dist
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback on the product path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    distributed = world_size > 1
    if distributed:
        dist.init_process_group(backend="nccl", device_id=dev)
    n_gpus = world_size
    N = args.envs
    K, W = max(args.steps, 1), max(args.warmup, 3)

    world = build_world(args.motions)
    env = build_gpu_env(world, N, rank * N, n_gpus * N, dev)
    init_gpu_state(env, rank * N)
    step_bytes, b_env = algorithmic_bytes(env)
    from parc_b200 import synth
    noise = synth.standin_noise(N, env._J, env._D, seed=5678 + rank * N, fail_fraction=args.fail_fraction).to(dev)   # simulator stand-in

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(ms):
        if not distributed:
            return ms
        tt = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    # ---- device-resident timing: K steps, per-launch events around the dominant kernel ---------------
    def final_stats():
        # final scalar reward / termination statistics: the only cross-rank exchange of the path
        st = torch.stack([env._reward_buf.sum().double(), (env._done_buf != 0).sum().double(),
                          env._motion_id_fail_rates.mean().double()])
        if distributed:
            dist.all_reduce(st, op=dist.ReduceOp.SUM)
        return st

    for _ in range(W - 1):
        env.sim_standin(noise)
        env.step()
    env.sim_standin(noise)
    env.post_physics_step()
    done_per_step = int(env._done_count.item())      # envs that finish (and are reset) in a steady-state step
    env.reset_done()
    final_stats()          # warm-up covers the statistics reduction too (first use loads its kernels / NCCL channels)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    from parc_b200 import _lib
    L = _lib.lib()
    km, lib_cs = env._kin_char_model._cstruct(), env._mlib.cstruct()
    stream = env._stream()
    barrier()
    t_start.record()
    host_t0 = time.perf_counter()
    for i in range(K):
        env.sim_standin(noise)                 # "physics": the character follows its reference with a fixed error
        ev[i][0].record()
        _lib.check(L.parc_b200_env_step(km, lib_cs, env._terrain_cs, env._cfg, env._bufs, _lib.STEP_ALL, None, 0, stream))
        ev[i][1].record()
        _lib.check(L.parc_b200_fail_rate_ema(env._cfg, env._bufs, env._mlib.num_motions(), _lib.ptr(env._cdf),
                                             _lib.ptr(env._ema_scratch), stream))
        env._step_count += 1
        env.reset_done()
    host_enqueue_ms = (time.perf_counter() - host_t0) * 1e3 / K
    stats = final_stats()
    t_end.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = max_over_ranks(t_start.elapsed_time(t_end))
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    kern_ms = max_over_ranks(kern_ms)
    value = n_gpus * N * K / (ms_total * 1e-3)

    # ---- end to end: host buffers in, host results out, every step -----------------------------------
    host_state = env.make_host_state()
    for k, v in env.sim_state_tensors().items():
        host_state[k].copy_(v)
    host_res = env.make_host_results()
    torch.cuda.synchronize(dev)
    h2d = sum(v.numel() * v.element_size() for v in host_state.values())
    d2h = sum(v.numel() * v.element_size() for v in host_res.values())
    Ke = K

    def time_e2e(one_step):
        for _ in range(3):
            one_step()
            torch.cuda.current_stream(dev).synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(Ke):
            one_step()
            torch.cuda.current_stream(dev).synchronize()   # the caller needs the results before the next physics step
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    def e2e_serial():
        env.upload_sim_state(host_state)       # this step's simulator state from pinned host memory, all of it first
        env.sim_standin(noise)                 # (its values are then re-centred on the reference: stationary workload)
        env.post_physics_step()
        env.download_results(host_res)         # reward + done back to the host
        env.reset_done()

    def e2e_chunked():
        # the public host-buffer call: parc_b200_env_step_host uploads range c+1 while range c is stepped
        env.step_host(host_state, host_res, num_chunks=args.chunks, standin_noise=noise)

    e2e_serial_ms = time_e2e(e2e_serial)
    e2e_ms = time_e2e(e2e_chunked)
    e2e_value = n_gpus * N * Ke / (e2e_ms * 1e-3)

    # ---- CPU baseline (rank 0, N=1 only) ------------------------------------------------------------
    cb = None
    if rank == 0 and n_gpus == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_baseline_block(world, args, 30, 5)

    if rank == 0:
        peak, peak_src = measured_peaks()
        achieved = step_bytes / (kern_ms * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "env_step_traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "host_enqueue_ms_per_step": host_enqueue_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "65,536 envs/GPU, 1,000-clip synthetic library of the repo humanoid, 32x32 local heightmap "
                                   "observations (BASELINE configs[1]); step = simulator stand-in (char := ref + fixed noise, FK) + post-physics "
                                   "pass + fail-rate EMA + device-side RSI reset + obs refresh of the reset envs",
                       "envs_per_gpu": N, "clips": args.motions, "frames": int(env._mlib._frame_root_pos.shape[0]),
                       "ray_points": int(env._ray_xy_points.shape[0]), "resets_per_step": done_per_step,
                       "standin_fail_fraction": args.fail_fraction, "obs_dim": env.get_obs_dim(),
                       "terrain_cells": list(env._terrain.hf.shape), "parallelism": "env-sharded x%d, no data-path collective" % n_gpus,
                       "l2": "per-step working set %.2f GB > 126 MB L2 (no flush needed)" % (step_bytes / 1e9)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / Ke, "chunks": args.chunks, "ms_per_step_unpipelined": e2e_serial_ms / Ke,
                    "note": "TrackingEnv.step_host -> parc_b200_env_step_host: every step uploads root/dof/rigid-body/contact state from "
                            "pinned host memory in row ranges (range c+1 in flight while range c is stepped), fail-rate EMA, D2H of "
                            "reward+done, RSI reset, host sync; the observation stays on the device for the policy, as in the reference"},
            "gpu_launches": LAUNCHES_PER_STEP * K,
            "roofline": {"bound": "hbm", "kernel": "env_step_kernel<512>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": step_bytes, "bytes_per_env": b_env,
                         "kernel_ms": kern_ms, "kernel_share_of_step": kern_ms / (ms_total / K)},
            "cpu_baseline": cb,
            "stats": {"reward_sum": float(stats[0]), "done_last_step": float(stats[1]), "mean_fail_rate": float(stats[2]) / n_gpus},
        }
        print(json.dumps(line))
    if distributed:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
