"""Samplers of the terrain node (chrono_wisc_b200/terrain.py) against the reference's definitions: host code, no GPU."""
import numpy as np

from chrono_wisc_b200 import terrain


def test_samplers_follow_the_reference_counts():
    # ChGridSampler: n = round(2 h / sep) divisions per axis, points on both ends (ChUtilsSamplers.h:506-521)
    p = terrain.grid_points((0.2, 0.0, 0.03), (0.19, 0.14, 0.02), 0.0204)
    n = [int(round(2 * h / 0.0204)) + 1 for h in (0.19, 0.14, 0.02)]
    assert p.shape[0] == n[0] * n[1] * n[2]
    assert np.allclose(p.min(0), [0.01, -0.14, 0.01]) and np.allclose(p.max(0), [0.39, 0.14, 0.05])
    # ChHCPSampler: no two points closer than the separation
    q = terrain.hcp_points((0.2, 0.0, 0.03), (0.1, 0.1, 0.02), 0.0204)
    d = np.linalg.norm(q[:, None] - q[None], axis=2) + np.eye(q.shape[0])
    assert d.min() > 0.0204 * (1 - 1e-9)
