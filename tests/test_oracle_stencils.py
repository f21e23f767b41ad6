"""Known-answer tests of the oracle's deriv.f90 restatement (CPU only).

The 6th-order stencils are exact for polynomials up to degree 6 (der: odd part up
to x^6 -> exact through degree 6; der2 exact through degree 7), and have the
textbook modified wavenumbers on sin(kx); the bidiagonal derij is exact for
x^a y^b with a+b <= 7 except its known leading error term."""
import numpy as np
import pytest

NG = 3


def padded_grid(n, dx):
    return (np.arange(-NG, n + NG) * dx)


def field_from(fx, n=12, d=(0.1, 0.13, 0.17)):
    x, y, z = padded_grid(n, d[0]), padded_grid(n, d[1]), padded_grid(n, d[2])
    Z, Y, X = np.meshgrid(z, y, x, indexing="ij")
    return np.ascontiguousarray(fx(X, Y, Z)), (X[NG:-NG, NG:-NG, NG:-NG], Y[NG:-NG, NG:-NG, NG:-NG], Z[NG:-NG, NG:-NG, NG:-NG]), d


@pytest.mark.parametrize("j", [1, 2, 3])
def test_der_exact_on_degree6_polynomials(orc, j):
    f, (X, Y, Z), d = field_from(lambda x, y, z: 1 + x + 2 * y ** 2 - z ** 3 + x ** 4 * y + y ** 5 - z ** 6 + x ** 6)
    got = orc.der(f, j, 1.0 / d[j - 1])
    exact = {1: 1 + 4 * X ** 3 * Y + 6 * X ** 5, 2: 4 * Y + X ** 4 + 5 * Y ** 4, 3: -3 * Z ** 2 - 6 * Z ** 5}[j]
    assert np.abs(got - exact).max() < 1e-11 * max(1.0, np.abs(exact).max())


@pytest.mark.parametrize("j", [1, 2, 3])
def test_der2_exact_on_degree7_polynomials(orc, j):
    f, (X, Y, Z), d = field_from(lambda x, y, z: x ** 7 - 3 * y ** 6 + z ** 5 + x ** 2 * y * z)
    got = orc.der2(f, j, 1.0 / d[j - 1])
    exact = {1: 42 * X ** 5 + 2 * Y * Z, 2: -90 * Y ** 4 + 0 * X, 3: 20 * Z ** 3 + 0 * X}[j]
    assert np.abs(got - exact).max() < 1e-9 * max(1.0, np.abs(exact).max())


def test_der6_upwind_is_sixth_difference(orc):
    # der6(upwind) = (dx_1/60) * delta^6 f ; for f=x^6, delta^6 f = 720 dx^6  ->  12 dx^5
    f, (X, Y, Z), d = field_from(lambda x, y, z: x ** 6 + 0 * y)
    got = orc.der6(f, 1, 1.0 / d[0])
    assert np.allclose(got, 12 * d[0] ** 5, rtol=1e-8, atol=0)
    f, _, d = field_from(lambda x, y, z: x ** 5 + y ** 3)
    assert np.abs(orc.der6(f, 1, 1.0 / d[0])).max() < 1e-12


@pytest.mark.parametrize("ij", [(1, 2), (2, 3), (3, 1), (2, 1), (1, 3)])
def test_derij_bidiagonal(orc, ij):
    i, j = ij
    f, (X, Y, Z), d = field_from(lambda x, y, z: x * y + 2 * y * z - 3 * z * x + x ** 3 * y ** 2 + y ** 3 * z ** 3 + z ** 2 * x ** 4)
    got = orc.derij(f, i, j, 1.0 / d[i - 1], 1.0 / d[j - 1])
    key = tuple(sorted(ij))
    exact = {(1, 2): 1 + 6 * X ** 2 * Y, (2, 3): 2 + 9 * Y ** 2 * Z ** 2, (1, 3): -3 + 8 * Z * X ** 3}[key]
    assert np.abs(got - exact).max() < 1e-10 * max(1.0, np.abs(exact).max())


def test_der_modified_wavenumber(orc):
    # 6th-order centred first derivative of sin(kx): k_eff dx = (45 sin a - 9 sin 2a + sin 3a)/30
    n, L = 32, 2 * np.pi
    dx = L / n
    x = (np.arange(-NG, n + NG) + 0.5) * dx
    k = 5.0
    f = np.ascontiguousarray(np.broadcast_to(np.sin(k * x), (n + 6, n + 6, n + 6)))
    got = orc.der(f, 1, 1.0 / dx)[0, 0]
    a = k * dx
    keff = (45 * np.sin(a) - 9 * np.sin(2 * a) + np.sin(3 * a)) / (30 * dx)
    assert np.allclose(got, keff * np.cos(k * x[NG:-NG]), atol=1e-13)


@pytest.mark.parametrize("ij", [(1, 2), (2, 3), (3, 1)])
def test_derij_36point_is_exact_for_degree6_products(orc, ij):
    """lbidiagonal_derij=F (deriv.f90:1430-1557): product of two 6th-order first-derivative
    stencils, exact for x^a y^b with a, b <= 6, and equal to der_j(der_i f) by construction."""
    i, j = ij
    f, (X, Y, Z), d = field_from(lambda x, y, z: x ** 6 * y ** 5 + y ** 6 * z ** 4 - 2 * z ** 6 * x ** 3 + x * y * z)
    got = orc.derij_full(f, i, j, 1.0 / d[i - 1], 1.0 / d[j - 1])
    exact = {(1, 2): 30 * X ** 5 * Y ** 4 + Z, (2, 3): 24 * Y ** 5 * Z ** 3 + X, (3, 1): -36 * Z ** 5 * X ** 2 + Y}[ij]
    assert np.abs(got - exact).max() < 1e-10 * max(1.0, np.abs(exact).max())
    bid = orc.derij(f, i, j, 1.0 / d[i - 1], 1.0 / d[j - 1])
    assert np.abs(bid - exact).max() > 1e-8          # the bidiagonal scheme is a different discretisation
