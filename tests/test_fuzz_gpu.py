"""A slice of scripts/fuzz_parity.py in the suite: random states in regimes the fixtures do not visit (jumps, smooth,
power-of-two scalings towards the edges of the fast division's window, fields vanishing in patches, out-of-plane
fields around 1e-160, an explosion-front decay), all boundary pairs, ragged sizes 4..100 -- CUDA == oracle in every
bit after 1, 3 and 6 steps; also through uneven slab groups, with other physical parameters, and with the table EoS.
The full runs (5600 cases) are summarised in profiles/r2/fuzz_parity_4000.log."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("first", [5000, 5025, 5050])
def test_mhd_fuzz(first):
    import fuzz_parity
    assert all([fuzz_parity.mhd_case(seed) for seed in range(first, first + 25)])


@pytest.mark.parametrize("first", [5000, 5020, 5040])
def test_euler_fuzz(first):
    import fuzz_parity
    assert all([fuzz_parity.euler_case(seed) for seed in range(first, first + 20)])


def test_fuzz_through_slab_groups():
    import fuzz_parity
    assert all([fuzz_parity.mhd_case(seed, slabs=True) for seed in range(6000, 6030)])
    assert all([fuzz_parity.euler_case(seed, slabs=True) for seed in range(6000, 6030)])


def test_fuzz_other_domains_courant_numbers_gammas_and_limiter_weights():
    import fuzz_parity
    assert all([fuzz_parity.mhd_case(seed, params=True) for seed in range(6100, 6120)])
    assert all([fuzz_parity.euler_case(seed, params=True) for seed in range(6100, 6120)])


def test_tabeos_fuzz():
    import fuzz_parity
    assert all([fuzz_parity.tabeos_case(seed) for seed in range(6200, 6220)])


def test_fuzz_initial_condition_generators():
    import fuzz_parity
    assert all([fuzz_parity.ic_case(seed) for seed in range(6300, 6320)])


def test_fuzz_long_single_call(monkeypatch):
    """19 steps in one call: CUDA graphs of 8 steps plus single steps of the device-resident loop"""
    import fuzz_parity
    monkeypatch.setattr(fuzz_parity, "SCHEDULE", (19,))
    assert all([fuzz_parity.mhd_case(seed) for seed in range(6400, 6415)])
    assert all([fuzz_parity.euler_case(seed) for seed in range(6400, 6415)])


def test_fuzz_contexts_uploaded_again_and_again(monkeypatch):
    import fuzz_parity
    cache = {}
    monkeypatch.setattr(fuzz_parity, "REUSE", cache)
    monkeypatch.setattr(fuzz_parity, "sizes", lambda rng, sd, small: (12 + 8 * (sd % 2), 20))
    try:
        assert all([fuzz_parity.mhd_case(seed) for seed in range(6500, 6530)])
        assert all([fuzz_parity.euler_case(seed) for seed in range(6500, 6530)])
    finally:
        for s in cache.values():
            del s.close          # the instance-level no-op: the class method closes for real
            s.close()


def test_fuzz_whole_loop_with_dumps_and_random_stop_times():
    import fuzz_parity
    assert all([fuzz_parity.run_case(seed) for seed in range(6600, 6630)])
