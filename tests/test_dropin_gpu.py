"""The compiled drop-in (oracle/pbrt_b200.cpp -> oracle/_ref/pbrt_b200): a .pbrt file parsed by the reference's own
parser, rendered on the GPU through libb200pt.so, written by the reference's own Film::WriteImage - against the same
file through the reference's CPU renderer (oracle/_ref/pbrt_ref).
  Level 2 (RenderB200 in place of RenderWavefront, wavefront/wavefront.cpp:14-70): rRMSE <= 1e-3 against the CPU
          PathIntegrator at the same sampler seed (the kernels consume the sampler dimensions in its order).
  Level 1 (B200Aggregate : WavefrontAggregate, wavefront/integrator.h:32-54, under the reference's own
          WavefrontPathIntegrator): the image of `pbrt_ref --wavefront` (CPUAggregate), to rounding of the film sums."""
import os
import re

import numpy as np
import pytest

from conftest import rrmse

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dropin(refrun):
    if not (refrun.b200_available() and refrun.render_available()):
        pytest.skip("oracle/_ref/pbrt_b200 not built (make -f oracle/ref.mk)")
    return refrun


def _scene(P, tmp_path, kind):
    pbrt, pfm = str(tmp_path / "s.pbrt"), str(tmp_path / "s.pfm")
    S = P.scenes
    if kind == "cornell":
        S.cornell_pbrt(pbrt, out=pfm, res=96, spp=16, maxdepth=5, lightsampler="power", seed=0)
    elif kind == "instanced":
        S.cornell_pbrt(pbrt, out=pfm, res=64, spp=8, maxdepth=4, lightsampler="bvh", seed=1, instanced=True)
    elif kind == "volpath":
        S.config5_pbrt(pbrt, out=pfm, res=(48, 32), spp=4, maxdepth=4, grid=16)
    return pbrt, pfm


@pytest.mark.parametrize("kind", ["cornell", "instanced"])
def test_render_b200_matches_pbrt_ref(P, dropin, tmp_path, kind):
    pbrt, pfm = _scene(P, tmp_path, kind)
    ref, _ = dropin.render(pbrt, pfm)
    os.unlink(pfm)
    img, _, log = dropin.render_b200(pbrt, pfm)
    assert "samples in" in log
    assert rrmse(img, ref) <= 1e-3, rrmse(img, ref)
    assert abs(img.mean() - ref.mean()) <= 1e-4 * ref.mean()
    # the library building the trees itself (same SAH decisions) instead of taking the host's BVHAggregate arrays
    os.unlink(pfm)
    img2, _, _ = dropin.render_b200(pbrt, pfm, extra=["--own-bvh"])
    assert rrmse(img2, ref) <= 1e-3


@pytest.mark.parametrize("kind", ["cornell", "instanced", "volpath"])
def test_wavefront_over_b200_aggregate(P, dropin, tmp_path, kind):
    pbrt, pfm = _scene(P, tmp_path, kind)
    ref = dropin.render_wavefront(pbrt, pfm)
    os.unlink(pfm)
    img, _, log = dropin.render_b200(pbrt, pfm, wavefront=True)
    m = re.search(r"B200Aggregate: (\d+) closest \+ (\d+) shadow rays on the GPU, (\d+) single-ray calls, (\d+) t mismatches", log)
    assert m, log
    closest, shadow, single, mismatches = (int(x) for x in m.groups())
    assert closest > 0 and mismatches == 0
    assert (shadow > 0) if kind != "volpath" else (single > 0)  # volpath: shadow rays walk through IntersectShadowTr
    assert rrmse(img, ref) <= 1e-5, rrmse(img, ref)


@pytest.mark.parametrize("film", ["gbuffer", "gbuffer_albedo"])
def test_render_b200_gbuffer_exr(P, dropin, tmp_path, film):
    """A .pbrt with the ISET film (`Film "gbuffer"`, EXR only) through the compiled drop-in: the reference's parser and
    scene objects, the GPU frame, the file written by b200pt_film_write_exr (device-side packing) - against the
    reference's own Film::GetImage of the same file (dumpscene_ref --render; the oracle build has no EXR writer)."""
    from conftest import read_exr
    pbrt, out = str(tmp_path / "g.pbrt"), str(tmp_path / "g.exr")
    P.scenes.cornell_pbrt(pbrt, out=out, res=48, spp=16, maxdepth=5, film=film, smooth="instanced")  # (_albedo: saveAlbedo)
    ref, names, _ = dropin.render_channels(pbrt)
    _, _, log = dropin.render_b200(pbrt, None)
    assert "samples in" in log
    ch, data, _ = read_exr(out)
    assert data == (0, 0, 47, 47) and set(ch) == set(names)
    for k, nm in enumerate(names):
        a, b = ch[nm].astype(np.float64), ref[..., k].astype(np.float64)
        if nm in ("MaterialId", "InstanceId"):
            assert np.array_equal(a, b), nm
        else:
            assert np.sqrt(np.mean((a - b) ** 2)) <= 1e-3 * max(np.mean(np.abs(b)), 1e-12), nm
