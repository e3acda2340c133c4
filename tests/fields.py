"""Named vector fields used by the parity tests and by tests/golden/make_golden.py.

Bytecode fixtures refer to callables by these names (``"fn:<name>"``)."""
import numpy as np


def adv_field(x, y):
    """Advection field of the reference's linear advection-diffusion example."""
    return np.stack(((3 * y - x), (2 - y + 0 * x)), axis=-1)


def swirl_field(x, y):
    return np.stack((np.sin(x) * np.cos(y) + 0.3, -np.cos(x) * np.sin(y) + 0.1 * x * y), axis=-1)


FIELDS = {"adv_field": adv_field, "swirl_field": swirl_field}


def deform(xi, eta):
    """Smooth non-affine deformation of the unit-square mesh."""
    return (xi + 0.1 * np.sin(np.pi * xi) * np.sin(np.pi * eta), eta - 0.1 * np.sin(np.pi * xi) * np.sin(np.pi * eta))


# the quads of the reference's tests/test_interior.py:135-142 plus two sheared ones
QUADS = np.array(
    [
        [(-1, -1), (1, -1), (1, 1), (-1, 1)],
        [(-2, -2), (2, -2), (2, 2), (-2, 2)],
        [(-0.1, -2), (0.1, -2), (0.1, 2), (-0.1, 2)],
        [(-2, -0.1), (2, -0.1), (2, 0.1), (-2, 0.1)],
        [(-1, -2), (2, 0), (1.75, 0.75), (1.0, 1.0)],
        [(-1, -1), (1, -1), (2, 1), (0, 1)],
        [(0, 0), (1, -1), (2, 0), (1, 1)],
    ],
    dtype=np.float64,
)
