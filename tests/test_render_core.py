"""Pieces of the sampler / spectrum code checked on their own against the reference's tables
(tests/golden/sobol_dims01.npy = SobolMatrices32[0..103], util/sobolmatrices.cpp:40-58)."""
import ctypes as C
import itertools
import os

import numpy as np

from conftest import ROOT


def _lib(rendercheck):
    L = C.CDLL(os.path.join(ROOT, "tests", "native", "_build", "librendercheck.so"))
    L.rendercheck_sobol_column.restype = C.c_uint32
    L.rendercheck_sobol_dim1.restype = C.c_uint32
    L.rendercheck_sobol_dim1.argtypes = [C.c_uint64]
    L.rendercheck_spectrum_at.restype = C.c_float
    L.rendercheck_spectrum_at.argtypes = [C.c_void_p, C.c_float]
    return L


def test_sobol_columns_match_reference_table(rendercheck):
    L = _lib(rendercheck)
    gold = np.load(os.path.join(ROOT, "tests", "golden", "sobol_dims01.npy"))
    mine = [L.rendercheck_sobol_column(d, j) for d in (0, 1) for j in range(52)]
    assert np.array_equal(np.array(mine, np.uint32), gold)
    src = "/root/reference/src/pbrt/util/sobolmatrices.cpp"
    if os.path.exists(src):  # the fixture itself against the source, where the reference is mounted
        import re
        body = open(src).read()
        body = body[body.index("{", body.index("SobolMatrices32[")) + 1:]
        vals = [int(m.group(0), 16) for m in itertools.islice(re.finditer(r"0x[0-9a-fA-F]+", body), 104)]
        assert np.array_equal(np.array(vals, np.uint32), gold)


def test_sobol_dim1_butterfly_is_the_matrix_product(rendercheck):
    L = _lib(rendercheck)
    gold = np.load(os.path.join(ROOT, "tests", "golden", "sobol_dims01.npy"))[52:]
    rng = np.random.default_rng(5)
    cases = [0, 1, 2, 3, (1 << 52) - 1, 1 << 51, 1 << 32, (1 << 32) - 1] + \
            [int(x) for x in rng.integers(0, 1 << 52, 2000, dtype=np.uint64)]
    for a in cases:
        v = 0
        for j in range(52):
            if (a >> j) & 1:
                v ^= int(gold[j])
        assert L.rendercheck_sobol_dim1(a) == v, hex(a)


def test_permutation_table(rendercheck):
    """samplers.h:300-326 lists the 24 permutations of {0,1,2,3} in this order."""
    L = _lib(rendercheck)
    ref = [(0, 1, 2, 3), (0, 1, 3, 2), (0, 2, 1, 3), (0, 2, 3, 1), (0, 3, 2, 1), (0, 3, 1, 2), (1, 0, 2, 3),
           (1, 0, 3, 2), (1, 2, 0, 3), (1, 2, 3, 0), (1, 3, 2, 0), (1, 3, 0, 2), (2, 1, 0, 3), (2, 1, 3, 0),
           (2, 0, 1, 3), (2, 0, 3, 1), (2, 3, 0, 1), (2, 3, 1, 0), (3, 1, 2, 0), (3, 1, 0, 2), (3, 2, 1, 0),
           (3, 2, 0, 1), (3, 0, 2, 1), (3, 0, 1, 2)]
    assert sorted(ref) == sorted(itertools.permutations(range(4)))
    for p in range(24):
        assert tuple(L.rendercheck_permute(p, d) for d in range(4)) == ref[p]


def test_nearest_spectrum_lookup_is_lround(P, rendercheck):
    """DenselySampledSpectrum::Sample (util/spectrum.h:420-430): values[lround(lambda) - lambda_min]."""
    L = _lib(rendercheck)
    s = P.capi.Spectrum()
    s.interp = 0
    for i in range(311):
        s.values[i] = float(i)
    rng = np.random.default_rng(3)
    lam = np.concatenate([rng.uniform(395, 705, 5000).astype(np.float32),
                          np.arange(395, 706, dtype=np.float32) + np.float32(0.5),
                          np.nextafter(np.arange(395, 706, dtype=np.float32) + np.float32(0.5), np.float32(0)),
                          np.array([394.4, 394.6, 705.4, 705.6, 0, 1000], np.float32)])
    for x in lam:
        off = int(np.floor(np.float64(x) + 0.5)) - 395  # lround: half away from zero, x > 0
        want = float(off) if 0 <= off < 311 else 0.0
        assert L.rendercheck_spectrum_at(C.byref(s), float(x)) == want, x


import pytest


@pytest.mark.parametrize("name", ["cornell_manylights_r64_s16", "cornell_infinite_r64_s16"])
def test_light_bvh_sample_and_pmf_match_reference(P, rendercheck, name):
    """BVHLightSampler::Sample / PMF (lightsamplers.h:266-358) at 768 random reference points against the
    answers the reference gave when the fixture was dumped (oracle/dumpscene_ref.cpp): the light chosen and
    both probabilities, bit for bit.  Then the properties the reference's own tests pin
    (lightsamplers_test.cpp:24-352): Sample's probability is PMF(light) of the light it returned."""
    rsc = P.RenderScene.load(os.path.join(ROOT, "tests", "golden", name + ".b2rs"))
    d = rsc.d
    keep = []
    sd = P.render.scene_desc(rsc.geometry, None, keep)
    rd = rsc.desc(keep)
    L = _lib(rendercheck)
    pin = np.ascontiguousarray(d["lightbvh_probe_in"], np.float32)
    query = np.ascontiguousarray(d["lightbvh_probe_light"][:, 1], np.int32)
    n = len(pin)
    light = np.zeros(n, np.int32)
    pmf = np.zeros((n, 2), np.float32)
    rc = L.rendercheck_lightbvh(C.byref(sd), C.byref(rd), pin.ctypes.data_as(C.c_void_p),
                                query.ctypes.data_as(C.c_void_p), n, light.ctypes.data_as(C.c_void_p),
                                pmf.ctypes.data_as(C.c_void_p))
    assert rc == 0
    assert np.array_equal(light, d["lightbvh_probe_light"][:, 0])
    assert np.array_equal(pmf.view(np.uint32), np.ascontiguousarray(d["lightbvh_probe_pmf"], np.float32).view(np.uint32))
    hit = light >= 0
    assert hit.sum() > n // 2
    # for a light of the tree Sample's probability is PMF(light) up to the order of the products
    # (infinite lights: PMF is exact)
    np.testing.assert_allclose(pmf[hit, 0], pmf[hit, 1], rtol=2e-6)
    nl = len(d["lights_i"])
    assert len(set(light[hit].tolist())) > min(nl, 20) // 2  # the descent reaches many different lights
