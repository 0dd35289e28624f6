"""parc_b200_fail_rate_ema_noalloc (include/parc_b200_env.h): the allocation-free variant of the fail-rate EMA, whose scratch
is a file-scope device array inside the library. Not a reference function;
checked against the existing entry point parc_b200_fail_rate_ema, which the other parity tests pin to the reference.
Collected last on purpose."""
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu


def _env_with_finished_envs(seed):
    sc = H.Scenario(num_envs=1003, num_motions=7, ray="grid", seed=seed, wrap_fraction=0.2, max_frames=90)
    oenv = sc.oracle_env()
    H.randomize_oracle_state(oenv, seed=seed + 2, fail_fraction=0.3)
    penv = sc.product_env("cuda:0")
    H.copy_oracle_state_to_product(oenv, penv)
    return penv


def _ema_both_ways(penv):
    """One fused step (fills done_list / done_code), then the EMA through both entry points from the same fail rates."""
    from parc_b200 import _lib
    L = _lib.lib()
    M = penv._mlib.num_motions()
    penv.post_physics_step(update_fail_rates=False)
    n_done = int(penv._done_count.item())
    assert n_done > 0
    rates0 = penv._motion_id_fail_rates.clone()
    # existing entry point, caller-provided scratch
    penv._cdf.zero_()
    _lib.check(L.parc_b200_fail_rate_ema(penv._cfg, penv._bufs, M, _lib.ptr(penv._cdf), _lib.ptr(penv._ema_scratch), penv._stream()))
    rates_a, cdf_a = penv._motion_id_fail_rates.clone(), penv._cdf.clone()
    # allocation-free entry point, same starting rates; outputs pre-allocated BEFORE the call
    penv._motion_id_fail_rates.copy_(rates0)
    penv._cdf.zero_()
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    _lib.check(L.parc_b200_fail_rate_ema_noalloc(penv._cfg, penv._bufs, M, _lib.ptr(penv._cdf), penv._stream()))
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before       # nothing went through torch's allocator
    torch.cuda.synchronize()
    rates_b, cdf_b = penv._motion_id_fail_rates.clone(), penv._cdf.clone()
    assert not torch.equal(rates_a, rates0), "the step must have changed some fail rate"
    # same kernel, same inputs: identical bits (the tolerance of the EMA parity tests is rtol 1e-6; equality is stricter)
    H.assert_close(rates_b.cpu(), rates_a.cpu(), rtol=1e-6, atol=0, what="fail rates")
    H.assert_close(cdf_b.cpu(), cdf_a.cpu(), rtol=1e-6, atol=0, what="rsi cdf")
    assert torch.equal(rates_b, rates_a) and torch.equal(cdf_b, cdf_a)
    return n_done


def test_noalloc_fail_rate_ema_equals_the_existing_entry_point_twice_in_a_row():
    penv = _env_with_finished_envs(seed=11)
    n1 = _ema_both_ways(penv)
    # second call in a row on the same env (its failing envs fail again): the built-in scratch still holds the first call's contents
    n2 = _ema_both_ways(penv)
    assert n1 > 0 and n2 > 0
    # and a different env (other sizes) right after, sharing the same array
    penv2 = _env_with_finished_envs(seed=23)
    _ema_both_ways(penv2)


def test_noalloc_fail_rate_ema_rejects_sizes_beyond_its_scratch():
    from parc_b200 import _lib
    penv = _env_with_finished_envs(seed=5)
    too_many = _lib.EMA_NOALLOC_MAX_INTS // 2 + 1          # 2 * M alone exceeds the array
    with pytest.raises(AssertionError, match="PARC_EMA_NOALLOC_MAX_INTS"):
        _lib.check(_lib.lib().parc_b200_fail_rate_ema_noalloc(penv._cfg, penv._bufs, too_many, None, penv._stream()))
