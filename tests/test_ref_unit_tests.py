"""The reference's own gtest suites for the hot path, run against the oracle build of the reference's sources
(oracle/ref.mk `test` -> oracle/_ref/pbrt_ref_test): the check that what the goldens were generated
with behaves as the reference's tests expect.  Skipped where the binary was not built (it is not part of
`make -f oracle/ref.mk all`; oracle/run_ref_tests.sh builds and runs the whole suite: 252 of 253 pass, the one
failure is RGBColorSpace.sRGB, 6.29e-3 against the fork's own 6e-3 tolerance for its 395-705 nm range)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "pbrt_ref_test")

# suites that exercise the functions on the path: shape intersection and sampling, bounds, the samplers
# and their low-discrepancy machinery, alias / piecewise tables, light sampling, phase function, float intervals
HOT_PATH = ("Triangle*:Bounds3*:Bounds2*:FloatInterval*:FloatingPoint*:ZSobolSampler*:Sobol*:LowDiscrepancy*:RNG*:"
            "PermutationElement*:AliasTable*:PiecewiseConstant1D*:PiecewiseConstant2D*:SampleDiscrete*:Sampling*:"
            "BVHLightSampling*:SpotLight*:DirectionCone*:HenyeyGreenstein*:Spectrum*:Hash*:Math*:Vector*:"
            "SphericalTriangleArea*:EqualArea*:Filter*")


@pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/pbrt_ref_test not built (make -f oracle/ref.mk test)")
def test_reference_gtests_for_the_path(tmp_path):
    r = subprocess.run([BIN, "--nthreads", "8", "--test-filter", HOT_PATH], cwd=tmp_path, capture_output=True,
                       text=True, timeout=900)
    tail = "\n".join(r.stdout.splitlines()[-15:])
    assert r.returncode == 0, tail
    passed = [ln for ln in r.stdout.splitlines() if ln.startswith("[  PASSED  ]")]
    assert passed and int(passed[0].split()[3]) >= 100, tail
