"""b200pt_ply_read (PLY ingestion; the reference's TriQuadMesh::ReadPLY, util/mesh.cpp:322-423): the three
encodings, the optional vertex properties under the reference's naming conventions, quads in bilinear-patch order, faces
of other lengths passed over, unknown properties and elements skipped, the reference's error cases - and the same file
through the reference's own `Shape "plymesh"` (oracle/_ref/dumpscene_ref): indices equal, vertices equal up to the
render-space translation."""
import os
import struct
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mesh(rng, nv=40, nt=30, nq=12):
    p = rng.uniform(-2, 2, (nv, 3)).astype(np.float32)
    n = rng.normal(size=(nv, 3)).astype(np.float32)
    uv = rng.uniform(0, 1, (nv, 2)).astype(np.float32)
    tri = rng.integers(0, nv, (nt, 3)).astype(np.int32)
    quad = rng.integers(0, nv, (nq, 4)).astype(np.int32)
    return p, n, uv, tri, quad


def _write(path, fmt, p, n, uv, tri, quad, uvnames=("u", "v"), extra=True, face_indices=True):
    """A PLY with a confusing but legal layout: an unknown element in front, unknown scalar and list properties between
    the known ones, a 5-gon among the faces."""
    faces = [list(t) for t in tri] + [list(q) for q in quad]
    faces.insert(3, [0, 1, 2, 3, 4])  # ignored by the reader (neither triangle nor quad)
    hdr = ["ply", f"format {fmt} 1.0", "comment made by tests/test_ply_read.py"]
    if extra:
        hdr += ["element material 2", "property uchar red", "property list uchar float weights"]
    hdr += [f"element vertex {len(p)}", "property float x", "property float y", "property double confidence", "property float z"]
    if n is not None:
        hdr += ["property float nx", "property float ny", "property float nz"]
    if uv is not None:
        hdr += [f"property float {uvnames[0]}", f"property float {uvnames[1]}"]
    hdr += [f"element face {len(faces)}", "property list uchar int vertex_indices"]
    if face_indices:
        hdr += ["property int face_indices"]
    hdr += ["end_header"]
    with open(path, "wb") as f:
        f.write(("\n".join(hdr) + "\n").encode())
        if fmt == "ascii":
            out = []
            if extra:
                out += ["7 2 0.5 0.25", "9 0"]
            for i in range(len(p)):
                row = [p[i, 0], p[i, 1], 0.75, p[i, 2]] + (list(n[i]) if n is not None else []) + (list(uv[i]) if uv is not None else [])
                out.append(" ".join(repr(float(x)) for x in row))
            for k, fc in enumerate(faces):
                out.append(" ".join(map(str, [len(fc)] + fc + ([100 + k] if face_indices else []))))
            f.write(("\n".join(out) + "\n").encode())
        else:
            e = "<" if fmt == "binary_little_endian" else ">"
            if extra:
                f.write(struct.pack(e + "BB2f", 7, 2, 0.5, 0.25) + struct.pack(e + "BB", 9, 0))
            for i in range(len(p)):
                f.write(struct.pack(e + "2fdf", p[i, 0], p[i, 1], 0.75, p[i, 2]))
                if n is not None:
                    f.write(struct.pack(e + "3f", *n[i]))
                if uv is not None:
                    f.write(struct.pack(e + "2f", *uv[i]))
            for k, fc in enumerate(faces):
                f.write(struct.pack(e + "B%di" % len(fc), len(fc), *fc))
                if face_indices:
                    f.write(struct.pack(e + "i", 100 + k))


@pytest.mark.parametrize("fmt", ["ascii", "binary_little_endian", "binary_big_endian"])
@pytest.mark.parametrize("uvnames", [("u", "v"), ("s", "t"), ("texture_u", "texture_v"), ("texture_s", "texture_t")])
def test_ply_read_round_trip(P, tmp_path, fmt, uvnames):
    p, n, uv, tri, quad = _mesh(np.random.default_rng(5))
    path = tmp_path / "m.ply"
    _write(path, fmt, p, n, uv, tri, quad, uvnames=uvnames)
    m = P.capi.read_ply(path)
    np.testing.assert_array_equal(m["p"], p)
    np.testing.assert_array_equal(m["n"], n)
    np.testing.assert_array_equal(m["uv"], uv)
    np.testing.assert_array_equal(m["tri"], tri)
    np.testing.assert_array_equal(m["quad"], quad[:, [0, 1, 3, 2]])  # as a bilinear patch (util/mesh.cpp:299-305)
    assert len(m["face_indices"]) == len(tri) + len(quad) + 1 and m["face_indices"][0] == 100


def test_ply_read_optional_parts_and_errors(P, tmp_path):
    p, n, uv, tri, quad = _mesh(np.random.default_rng(6))
    path = tmp_path / "m.ply"
    _write(path, "binary_little_endian", p, None, None, tri, quad[:0], extra=False, face_indices=False)
    m = P.capi.read_ply(path)
    assert m["n"] is None and m["uv"] is None and m["quad"] is None and m["face_indices"] is None
    np.testing.assert_array_equal(m["tri"], tri)
    # scenes.write_ply (what the config-3 generator hands to the reference) reads back
    sc = P.scenes.random_soup(200, seed=3)
    P.scenes.write_ply(path, sc.meshes[0])
    m = P.capi.read_ply(path)
    np.testing.assert_array_equal(m["p"], sc.meshes[0].p.astype(np.float32))
    np.testing.assert_array_equal(m["tri"], sc.meshes[0].indices.reshape(-1, 3))
    # the reference's failure cases, with its words
    bad = tri.copy()
    bad[4, 1] = len(p)
    _write(path, "ascii", p, n, uv, bad, quad)
    with pytest.raises(P.capi.B200PTError, match="out of bounds"):
        P.capi.read_ply(path)
    with open(path, "wb") as f:
        f.write(b"ply\nformat ascii 1.0\nelement vertex 1\nproperty float x\nproperty float y\nelement face 1\n"
                b"property list uchar int vertex_indices\nend_header\n0 0\n3 0 0 0\n")
    with pytest.raises(P.capi.B200PTError, match="Vertex coordinate property not found"):
        P.capi.read_ply(path)
    with open(path, "wb") as f:
        f.write(b"\x1f\x8b\x08\x00junk")
    with pytest.raises(P.capi.B200PTError, match="gzip"):
        P.capi.read_ply(path)
    _write(path, "binary_little_endian", p, n, uv, tri, quad)
    data = open(path, "rb").read()
    open(path, "wb").write(data[:len(data) - 40])
    with pytest.raises(P.capi.B200PTError, match="end of file"):
        P.capi.read_ply(path)
    with pytest.raises(P.capi.B200PTError, match="cannot open"):
        P.capi.read_ply(tmp_path / "nothing.ply")


def test_ply_read_agrees_with_the_reference_reader(P, tmp_path):
    """The file through `Shape "plymesh"` of the compiled reference (its rply-based reader), dumped by dumpscene_ref."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    if not refrun.render_available():
        pytest.skip("oracle/_ref not built")
    p, n, uv, tri, quad = _mesh(np.random.default_rng(7), nv=60, nt=50, nq=0)
    ply = tmp_path / "m.ply"
    # (no face_indices here: with a face of another length in the file the reference CHECK-fails on their count, mesh.cpp:66)
    _write(ply, "binary_big_endian", p, n, uv, tri, quad, face_indices=False)
    pbrt = tmp_path / "s.pbrt"
    pbrt.write_text('LookAt 0 0 -9  0 0 0  0 1 0\nCamera "perspective" "float fov" 40\nSampler "zsobol" "integer pixelsamples" 1\n'
                    'Integrator "path" "integer maxdepth" 1\nPixelFilter "box"\nFilm "rgb" "string filename" "o.pfm" '
                    '"integer xresolution" 8 "integer yresolution" 8\nWorldBegin\n'
                    'AttributeBegin\n  AreaLightSource "diffuse" "rgb L" [1 1 1]\n  Material "diffuse"\n'
                    f'  Shape "plymesh" "string filename" "{ply}"\nAttributeEnd\n')
    d = P.scenes.load_b2rs(refrun.dump_scene(str(pbrt), str(tmp_path / "s.b2rs")))
    m = P.capi.read_ply(ply)
    np.testing.assert_array_equal(d["indices"].reshape(-1, 3), m["tri"])
    np.testing.assert_allclose(d["P"] - d["P"][0], m["p"] - m["p"][0], atol=2e-6)  # render space = world - eye
    np.testing.assert_allclose(d["corner_uv"].reshape(-1, 3, 2), m["uv"][m["tri"]], atol=0)
