"""Known-answer vectors restated from the reference's own test-suite (file:line
given per item, paths relative to the reference checkout).  Shared by the
oracle tests (CPU) and the CUDA parity tests (GPU)."""
import numpy as np

Node = np.dtype([('parent', 'uint32'), ('right_edge', 'uint32'), ('data', 'uint32', 2)])
NO_NODE = 0xFFFFFFFF

# tests/test_collision.py:251-265 (test_codes), also :302-311 (test_traverse)
SIX_COORDS = np.array([[0.0, 1.0, 3.0], [0.0, 1.0, 3.0], [4.0, 1.0, 8.0],
                       [-4.0, -6.0, 3.0], [-5.0, 0.0, -1.0], [-5.0, 0.5, -0.5]])
SIX_CODES = np.array([862940378, 862940378, 1073741823, 20332620, 302580864, 306295426], dtype='uint32')
SIX_PAIRS = {(0, 1), (4, 5)}           # tests/test_collision.py:421-422

# tests/test_collision.py:81-83, :109-123 (test_generate_bvh; Karras Fig. 3)
FIG3_CODES = np.array([0b00001, 0b00010, 0b00100, 0b00101, 0b10011, 0b11000, 0b11001, 0b11110], dtype='uint32')


def fig3_expected():
    leaf = 7
    nodes = np.array([(NO_NODE, 7, [3, 4]), (3, 1, [leaf + 0, leaf + 1]), (3, 3, [leaf + 2, leaf + 3]),
                      (0, 3, [1, 2]), (0, 7, [leaf + 4, 5]), (4, 7, [6, leaf + 7]),
                      (5, 6, [leaf + 5, leaf + 6])], dtype=Node)
    leaf_parents = np.array([1, 1, 2, 2, 4, 6, 6, 5])
    return nodes, leaf_parents


# tests/test_collision.py:135-136, :166-178 (test_generate_odd_bvh)
ODD_CODES = FIG3_CODES[:7].copy()


def odd_expected():
    leaf = 6
    nodes = np.array([(NO_NODE, 6, [3, 4]), (3, 1, [leaf + 0, leaf + 1]), (3, 3, [leaf + 2, leaf + 3]),
                      (0, 3, [1, 2]), (0, 6, [leaf + 4, 5]), (4, 6, [leaf + 5, leaf + 6])], dtype=Node)
    leaf_parents = np.array([1, 1, 2, 2, 4, 5, 5])
    return nodes, leaf_parents


# tests/test_collision.py:428-448 (test_problem_codes)
PROBLEM_CODES = np.array([
    0b00000000000000000000000000000000, 0b00000000000000000000000000000000,
    0b00000110110000110100000100000010, 0b00001001001001001001001001001001,
    0b00001001001001001001001001001001, 0b00010010010010010010010010010010,
    0b00010010010010010010010010010010, 0b00010010011010010010011011011010,
    0b00011001001011001001011001001011, 0b00011011011011011011011011011011,
    0b00100100010000100010110100010110, 0b00100100100100100100100100100100,
    0b00100100100101101101100101100100, 0b00101001101001101101101101101001,
    0b00101101101101101101101101101101, 0b00110110110110110110110110110110,
    0b00110110110110110110110110110110, 0b00110110110110110110110110110110,
    0b00111111111111111111111111111111, 0b00111111111111111111111111111111,
    0b00111111111111111111111111111111], dtype='uint32')

# tests/test_collision.py:185-195, :241-247 (test_compute_bounds)
REFIT_COORDS = np.array([[0.0, 1.0, 3.0], [4.0, 1.0, 8.0], [-4.0, -6.0, 3.0], [-5.0, 0.0, -1.0]])


def refit_nodes():
    leaf = 3
    return np.array([(NO_NODE, 3, [leaf + 0, 1]), (0, 3, [leaf + 3, 2]), (1, 2, [leaf + 1, leaf + 2]),
                     (0, 0, [2, NO_NODE]), (2, 1, [0, NO_NODE]), (2, 2, [1, NO_NODE]),
                     (1, 3, [3, NO_NODE])], dtype=Node)


REFIT_EXPECTED = np.array([[[-6.0, -7.0, -2.0], [5.0, 2.0, 9.0]], [[-6.0, -1.0, -2.0], [5.0, 2.0, 9.0]],
                           [[-1.0, 0.0, 2.0], [5.0, 2.0, 9.0]], [[-5.0, -7.0, 2.0], [-3.0, -5.0, 4.0]],
                           [[-1.0, 0.0, 2.0], [1.0, 2.0, 4.0]], [[3.0, 0.0, 7.0], [5.0, 2.0, 9.0]],
                           [[-6.0, -1.0, -2.0], [-4.0, 1.0, 0.0]]])

# tests/test_offset.py:31-47 and tests/test_offset_py.py:23-62
OFFSET_VALUES = np.array([0, 0, 1, 1, 1, 1, 1, 3, 3, 3, 4, 5, 5], dtype='uint32')
OFFSET_EXPECTED = np.array([0, 2, 7, 7, 10, 11, 13], dtype='uint32')


def slots(xyz, dtype):
    """(n, 3) -> (n, 4) slots with a poisoned pad lane (the reference leaves it
    uninitialised: tests/test_collision_py.py:63-72)."""
    out = np.full((len(xyz), 4), np.nan, dtype)
    out[:, :3] = xyz
    return out


def random_spheres(n, seed=4, dtype='float32', pad=0.0):
    """tests/test_collision_py.py:100-118: coords U[0,1)^3, radii U(0, 1/sqrt(n))."""
    rng = np.random.RandomState(seed)
    c = np.full((n, 4), pad, dtype)
    c[:, :3] = rng.random_sample((n, 3))
    r = rng.uniform(0, 1 / np.sqrt(n), n).astype(dtype)
    return c, r


def canon(pairs):
    """Row-sorted pair set (the reference's comparison form,
    tests/test_collision_py.py:148-150)."""
    return set(map(tuple, np.sort(np.asarray(pairs), axis=1).tolist()))


def lexsorted(pairs):
    p = np.asarray(pairs).reshape(-1, 2)
    return p[np.lexsort((p[:, 1], p[:, 0]))]


AWKWARD_SHAPES = ("uniform", "clusters", "duplicates", "diagonal", "lattice", "plane")


def awkward_scene(shape, seed, dtype='float32'):
    """Randomised small scenes of the kinds that break LBVH builders: runs of identical Morton codes
    (duplicates, lattices), collinear and coplanar centres, radii from zero to scene-sized.  Every axis
    keeps a non-zero extent (two corner spheres): a zero extent is 0/0 in the reference (undefined)."""
    rs = np.random.RandomState(1000 * AWKWARD_SHAPES.index(shape) + seed)
    n = int(rs.randint(2, 260))
    xyz = rs.random_sample((n, 3))
    if shape == "clusters":
        xyz = rs.random_sample((4, 3))[rs.randint(0, 4, n)] + rs.normal(0, 0.005, (n, 3))
    elif shape == "duplicates":
        xyz = xyz[rs.randint(0, max(n // 8, 1), n)]
    elif shape == "diagonal":
        xyz = np.repeat(rs.random_sample((n, 1)), 3, axis=1)
    elif shape == "lattice":
        xyz = rs.randint(0, 4, (n, 3)) / 3.0
    elif shape == "plane":
        xyz[:, 2] = 0.5
    c = np.zeros((n, 4), dtype)
    c[:, :3] = xyz
    c[0, :3], c[n - 1, :3] = 0.0, 1.0
    scale = (0.0, 0.01, 1 / np.sqrt(n), 0.6)[seed % 4]
    r = (rs.random_sample(n) * scale).astype(dtype)
    return c, r
