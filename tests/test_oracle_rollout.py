"""The rollout oracle against the golden vectors of the unmodified reference ParticleModel
(tests/golden/rollout_maze.npz, oracle/make_golden_rollout.py)."""
import numpy as np
import pytest

from oracle import rollout_oracle as ro
from tests.util import load_golden, rel_max


def golden_case(name):
    g = load_golden("rollout_maze")
    grid = np.unpackbits(g["grid_bits"])[: int(np.prod(g["grid_shape"]))].reshape(tuple(g["grid_shape"]))
    return g, grid, dict(dt=float(g["dt"]), max_speed=float(g["max_speed"]), cell_size=float(g["cell_size"]),
                         c_offset=tuple(g["c_offset"]), can_crash=True)


@pytest.mark.parametrize("name", ["free", "wild"])
def test_rollout_oracle_is_bit_identical_to_the_reference_model(name):
    g, grid, kw = golden_case(name)
    states, flags = ro.particle_rollout_np(g["init_state"], g[f"{name}_actions"], g[f"{name}_mass"], grid=grid, **kw)
    assert np.array_equal(states, g[f"{name}_states"])
    if name == "wild":
        assert flags[..., 0].any() and flags[..., 1:3].any()  # crashes and speed clamps are exercised
    ga = ro.particle_rollout_bwd_np(g[f"{name}_w"], flags, g[f"{name}_mass"], kw["dt"])
    assert rel_max(ga, g[f"{name}_grad_actions"]) < 1e-6
