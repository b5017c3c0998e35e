"""tests/golden/*.npz (answers of the compiled reference, made by tests/golden/make_golden.py)
against (a) the reference re-run now, when oracle/_ref is present, so the fixtures cannot go
stale, and (b) the host build of the device arithmetic, which needs no oracle binary."""
import importlib.util
import os

import numpy as np
import pytest

from conftest import ROOT, hit_mismatches
from test_hostcheck import run_hostcheck

_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
G = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(G)
NAMES = ["soup20k_primary", "soup20k_adversarial", "cornell_camera", "cornell_adversarial", "sheets_adversarial",
         "instanced_primary", "instanced_adversarial", "patches_primary", "patches_adversarial",
         "alpha_primary", "alpha_adversarial"]


@pytest.mark.parametrize("name", NAMES)
def test_fixture_matches_reference_rerun(P, refrun, name):
    sc, rays = G.cases(P.scenes)[name]
    gold = G.load(name)
    rs = refrun.RefScene(sc)
    hits, _ = rs.closest(rays)
    occ, _ = rs.any(rays)
    assert np.array_equal(G.compact(hits).view("u1"), gold["hits"].view("u1"))
    assert np.array_equal(occ, gold["occ"])


@pytest.mark.parametrize("name", NAMES)
def test_host_arithmetic_matches_fixture(P, hostcheck_bin, name, tmp_path):
    sc, rays = G.cases(P.scenes)[name]
    gold = G.load(name)

    class RS:  # scene file without needing the oracle binary
        dir = str(tmp_path)
        path = os.path.join(str(tmp_path), "scene.b2sc")
    sc.write(RS.path)
    mine, occ = run_hostcheck(hostcheck_bin, P, RS, rays)
    assert hit_mismatches(mine, gold["hits"], sc).sum() == 0
    assert np.array_equal(occ, gold["occ"])
    sh = P.scenes.shadow_rays(rays, gold["hits"], seed=35)
    mine, occ = run_hostcheck(hostcheck_bin, P, RS, sh)
    assert hit_mismatches(mine, gold["shadow_hits"], sc).sum() == 0
    assert np.array_equal(occ, gold["shadow_occ"])
