"""CPU: the slab-wise generator used by the 512^3 / 1024^3 multi-GPU runs yields ONE particle set whatever the partition."""
import numpy as np

import vrsynth


def test_make_slab_is_partition_invariant():
    n = 48 ** 3
    full = vrsynth.make_slab(n, 9, 0.0, vrsynth.BOX)
    assert full["n"] == n and np.array_equal(full["pid"], np.arange(1, n + 1))   # sorted by pid, every id once
    assert full["pos"].min() >= 0.0 and full["pos"].max() < vrsynth.BOX
    for world in (2, 8):
        parts = [vrsynth.make_slab(n, 9, vrsynth.BOX * r / world, vrsynth.BOX * (r + 1) / world) for r in range(world)]
        assert sum(p["n"] for p in parts) == n
        for r, p in enumerate(parts):
            x = p["pos"][:, 0]
            assert np.all((x >= vrsynth.BOX * r / world) & (x < vrsynth.BOX * (r + 1) / world))
            assert np.all(np.diff(p["pid"]) > 0)                                  # array order == pid order on every rank
        pid = np.concatenate([p["pid"] for p in parts])
        o = np.argsort(pid)
        assert np.array_equal(pid[o], full["pid"])
        assert np.array_equal(np.concatenate([p["pos"] for p in parts])[o], full["pos"])
        assert np.array_equal(np.concatenate([p["vel"] for p in parts])[o], full["vel"])
