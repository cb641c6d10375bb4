"""`_generate_positions` against the reference's golden vectors (src/simulation.rs:168-219):
bit-exact for the Python restatement (oracle) and for the C ABI (product)."""
from pathlib import Path

import numpy as np

import cr_mech_coli_b200 as crm
from oracle.generate_positions import generate_positions as oracle_generate

GOLD = Path(__file__).parent / "golden"


def test_oracle_matches_reference_golden():
    gold = np.load(GOLD / "generate_positions_seed1.npy")
    got = np.stack(oracle_generate(4, 3.0, (100.0, 100.0), 2.5, rng_seed=1, dx=(0.0, 0.0), randomize_positions=0.1,
                                   n_vertices=8))
    assert got.dtype == np.float32
    assert np.array_equal(got, gold)


def test_product_matches_reference_golden():
    gold = np.load(GOLD / "generate_positions_seed1.npy")
    got = np.stack(crm.generate_positions(4, crm.AgentSettings(), crm.Configuration(), rng_seed=1, dx=[0.0, 0.0],
                                          randomize_positions=0.1, n_vertices=8))
    assert np.array_equal(got, gold)


def test_product_matches_oracle_other_seeds():
    # non-square agent counts exercise the reservoir sampling (choose_multiple) draws
    for n, seed, jitter, nv in [(4, 0, 0.0, 8), (5, 3, 0.2, 8), (7, 11, 0.05, 6), (16, 2, 0.0, 4), (3, 99, 1.5, 8)]:
        cfg = crm.Configuration(domain_size=(400.0, 300.0))
        a = np.stack(crm.generate_positions(n, crm.AgentSettings(), cfg, rng_seed=seed, dx=[1.0, 2.0],
                                            randomize_positions=jitter, n_vertices=nv))
        b = np.stack(oracle_generate(n, 3.0, (400.0, 300.0), 2.5, rng_seed=seed, dx=(1.0, 2.0),
                                     randomize_positions=jitter, n_vertices=nv))
        assert np.array_equal(a, b), (n, seed)


def test_readme_fixture_is_seed0():
    gold = np.load(GOLD / "readme_positions_seed0.npy")
    agents = crm.generate_agents(4, crm.AgentSettings(), crm.Configuration())
    assert np.array_equal(np.stack([a.pos for a in agents]), gold)
    # rods stay inside their chunk and are straight with spring_length spacing
    d = np.linalg.norm(np.diff(gold, axis=1), axis=2)
    assert np.allclose(d, 3.0, atol=1e-4)


def test_too_small_domain_raises():
    import pytest

    with pytest.raises(ValueError):
        crm.generate_positions(100, crm.AgentSettings(), crm.Configuration())
