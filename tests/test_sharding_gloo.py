"""World-size-2 test of the N>1 path on CPU (gloo): envs split into contiguous blocks, no data-path
exchange, one sum of statistics at the end. The per-env results of a shard must be exactly the rows
of the single-process run (checked here with the oracle, on the GPU by
test_full_size_properties_65536_envs / bench.py --gpus N)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from parc_b200 import sharding
from tests import helpers as H

N_TOTAL, M = 45, 5


def _scenario():
    return H.Scenario(num_envs=N_TOTAL, num_motions=M, ray="cone", seed=21, wrap_fraction=0.2, max_frames=80)


def _run_rows(first, count):
    """Oracle step of global envs [first, first+count) cut out of the N_TOTAL scenario."""
    sc = _scenario()
    full = sc.oracle_env()
    H.randomize_oracle_state(full, seed=8)
    sub = H.Scenario.__new__(H.Scenario)
    sub.__dict__.update(sc.__dict__)
    sub.N, sub.env_offsets = count, sc.env_offsets[first:first + count]
    env = sub.oracle_env()
    sl = slice(first, first + count)
    for name in ("motion_ids", "motion_terrain_ids", "motion_time_offsets", "timestep_buf", "time_buf", "char_root_pos", "char_root_rot",
                 "char_root_vel", "char_root_ang_vel", "char_dof_pos", "char_dof_vel", "char_rigid_body_pos", "char_contact_forces"):
        getattr(env, name)[:] = getattr(full, name)[sl]
    env.post_physics_step()
    return env


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        first, count = sharding.shard_range(N_TOTAL, rank, world)
        env = _run_rows(first, count)
        stats = sharding.reduce_stats(sharding.local_stats(env.reward_buf, env.done_buf))
        code = torch.where(env.done_buf == 1, 1, torch.where(env.done_buf != 0, 2, 0)).to(torch.uint8)
        counts = sharding.motion_outcome_counts(code, env.motion_ids, M)
        dist.all_reduce(counts)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), first=first, count=count, obs=env.obs_buf.numpy(), reward=env.reward_buf.numpy(),
                 done=env.done_buf.numpy(), stats=stats.numpy(), counts=counts.numpy())
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_ranges_cover_all_envs_once():
    for total in (0, 1, 7, 45, 65536, 1048576):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (f0, c0), (f1, _) in zip(spans, spans[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
            for e in {0, total // 3, total - 1} - {-1}:
                if total:
                    r = sharding.owner_of(e, total, world)
                    f, c = spans[r]
                    assert f <= e < f + c


def test_two_rank_gloo_run_equals_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    whole = _run_rows(0, N_TOTAL)
    parts = [np.load(tmp_path / ("rank%d.npz" % r)) for r in range(world)]
    assert [int(p["first"]) for p in parts] == [0, 23] and [int(p["count"]) for p in parts] == [23, 22]
    # per-env outputs: a shard's rows are the same rows of the single-process run. torch's CPU kernels take a SIMD or a
    # scalar-tail path depending on where a row falls in the batch, so the CPU oracle itself is only reproducible to an
    # ulp across batch splits; flags are exact. (The CUDA path is bit-identical across shards: test_gpu_parity.py.)
    obs = np.concatenate([p["obs"] for p in parts])
    np.testing.assert_allclose(obs, whole.obs_buf.numpy(), rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(np.concatenate([p["reward"] for p in parts]), whole.reward_buf.numpy(), rtol=1e-5, atol=1e-6)
    assert np.array_equal(np.concatenate([p["done"] for p in parts]), whole.done_buf.numpy())
    # the one exchange: summed statistics are identical on both ranks and equal the single-process totals
    want = sharding.local_stats(whole.reward_buf, whole.done_buf).numpy()
    for p in parts:
        np.testing.assert_allclose(p["stats"], want, rtol=1e-6)
        assert np.array_equal(p["stats"], parts[0]["stats"])
        assert p["stats"][3] == N_TOTAL
    assert np.array_equal(parts[0]["counts"], parts[1]["counts"])
    code = torch.where(whole.done_buf == 1, 1, torch.where(whole.done_buf != 0, 2, 0)).to(torch.uint8)
    assert np.array_equal(parts[0]["counts"], sharding.motion_outcome_counts(code, whole.motion_ids, M).numpy())


def test_fail_rate_update_from_counts_matches_sequential_rule():
    """apply_outcome_counts == the reference's per-env loop (dm_env.py:648-660) with the passes of a
    motion applied before its fails, to fp32 rounding."""
    g = torch.Generator().manual_seed(3)
    rates = torch.rand(M, generator=g)
    counts = torch.randint(0, 6, (2, M), generator=g)
    want = rates.clone()
    for m in range(M):
        r = float(want[m])
        for _ in range(int(counts[1, m])):
            r = r * 0.99
        for _ in range(int(counts[0, m])):
            r = r * 0.99 + 0.01
        want[m] = r
    got = sharding.apply_outcome_counts(rates.clone(), counts)
    torch.testing.assert_close(got, want, rtol=2e-6, atol=1e-7)
