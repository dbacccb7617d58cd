"""Hand-made 3DFOF inputs for the corners of the domain (shared by the CPU pinning test against the compiled reference,
the serial emulation of the candidate-queue linker and the GPU parity tests): coincident particles, particles exactly on
the faces of the periodic box, groups of exactly minsize / minsize - 1 members, everything in one cell, fewer particles
than minsize, chains that close through two periodic faces at once."""
import numpy as np

L = 100.0
B = 1.0


def cases():
    """name -> (pos[n,3], period or None, minsize)."""
    rng = np.random.default_rng(17)
    out = {}
    # coincident particles (distance 0 < b) in three stacks, one of them split over the x = 0 / x = L face
    stack = np.repeat(np.array([[10.0, 20.0, 30.0], [50.0, 50.0, 50.0], [0.0, 70.0, 70.0]]), 5, axis=0)
    stack[12:15, 0] = L                       # same point as x = 0 in the periodic box
    out["coincident"] = (stack, L, 3)
    # particles exactly on faces, edges and corners; neighbours a little inside the opposite faces
    faces = np.array([[0.0, 0.0, 0.0], [L, L, L], [L - 0.4, 0.3, L], [0.3, L - 0.3, 0.0], [L, 0.0, 50.0], [0.0, L, 50.4],
                      [50.0, 0.0, L], [50.5, L, 0.0], [25.0, 25.0, 25.0]])
    out["faces"] = (faces, L, 2)
    # clumps of exactly minsize - 1, minsize and minsize + 1 members
    clumps = [c + 0.05 * rng.random((k, 3)) for c, k in ((10.0, 19), (30.0, 20), (50.0, 21), (70.0, 20))]
    out["minsize"] = (np.concatenate(clumps), L, 20)
    # everything inside one fine cell
    out["one_cell"] = (42.0 + 1e-3 * rng.random((64, 3)), L, 2)
    # fewer particles than minsize, and a single particle
    out["tiny"] = (np.array([[1.0, 1.0, 1.0], [1.2, 1.0, 1.0], [1.4, 1.0, 1.0]]), L, 20)
    out["single"] = (np.array([[5.0, 5.0, 5.0]]), L, 1)
    # a chain along the space diagonal that closes through all three periodic faces (spacing 0.9 b / sqrt(3) per axis)
    step = 0.9 * B / np.sqrt(3.0)
    t = (L - 3.0 + step * np.arange(12)) % L
    out["corner_chain"] = (np.column_stack([t, t, t]), L, 2)
    # distance exactly b along each axis (not linked: strict <) next to pairs one ulp closer (linked)
    exact = []
    for k in range(3):
        a = np.array([20.0, 20.0, 20.0]) + 10.0 * k
        bb = a.copy(); bb[k] += B
        c = a + 3.0
        d = c.copy(); d[k] = np.nextafter(c[k] + B, 0.0)
        exact += [a, bb, c, d]
    out["strict"] = (np.array(exact), L, 2)
    # the same diagonal chain in an open box (no wrap: it breaks where the coordinates jump)
    out["corner_chain_open"] = (np.column_stack([t, t, t]), None, 2)
    # KDTree::FOF applies the periodic images in its SPLIT nodes only (LeafNode::FOFSearchBallPeriodic, KDLeafNode.cxx:824-827,
    # just forwards to the open search), so an input of at most BucketSize particles -- a tree that is one leaf -- is
    # linked as if the box were open.  Every case therefore carries 24 isolated filler particles (mutual distances > 5 b,
    # > 5 b from every face and from the case's own particles), which makes the reference build at least one split node.
    g = np.arange(24)
    filler = np.column_stack([8.0 + 12.0 * (g % 4), 15.0 + 13.0 * ((g // 4) % 3), 88.0 - 9.0 * (g // 12)]) + 0.37
    for name, (pos, period, minsize) in list(out.items()):
        d = np.min(np.linalg.norm(pos[:, None, :] - filler[None, :, :], axis=2))
        assert d > 5 * B, (name, d)
        out[name] = (np.concatenate([pos, filler]), period, minsize)
    return out
