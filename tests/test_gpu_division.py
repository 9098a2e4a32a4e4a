"""The kernels divide a vec3 by a scalar with one shared reciprocal (csrc/trt_device.cuh div3_rn).
It must equal the IEEE-754 `/` operator bit for bit on every input, including zeros, denormals,
infinities and NaNs (those take the plain-division fallback)."""
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dist,n", [(0, 1 << 28), (1, 1 << 30), (2, 1 << 28)])
def test_div3_matches_ieee_division(pkg, dist, n):
    for seed in (1, 0xDEADBEEF):
        bad, first = pkg.api.selftest_div3(n, seed=seed, distribution=dist)
        assert bad == 0, f"distribution {dist}: {bad} mismatches, first (a, s, a/s, got) = {first}"
