// ply_read.cpp - b200pt_ply_read: a PLY mesh file into the arrays b200pt_scene_create takes, the way the reference's
// `Shape "plymesh"` reads one (TriQuadMesh::ReadPLY, util/mesh.cpp:322-423, through rply): vertex x y z, optional
// nx ny nz, optional (u, v) under the four naming conventions the reference accepts, faces of three or four vertices
// from the list property vertex_indices (other lengths are passed over; a quad comes out in bilinear-patch order
// 0 1 3 2, mesh.cpp:299-305), optional face_indices.  Everything else in the file is parsed far enough to be skipped.
// ascii, binary_little_endian and binary_big_endian; a gzip-compressed file is refused (the reference inflates it).
// Host only: no CUDA here, the arrays are plain malloc'ed memory the caller hands on to b200pt_scene_create.
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b200pt.h"

namespace {

enum Type { T_I8, T_U8, T_I16, T_U16, T_I32, T_U32, T_F32, T_F64, T_BAD };
const int kSize[] = {1, 1, 2, 2, 4, 4, 4, 8, 0};

Type TypeOf(const std::string &s) {
    if (s == "char" || s == "int8") return T_I8;
    if (s == "uchar" || s == "uint8") return T_U8;
    if (s == "short" || s == "int16") return T_I16;
    if (s == "ushort" || s == "uint16") return T_U16;
    if (s == "int" || s == "int32") return T_I32;
    if (s == "uint" || s == "uint32") return T_U32;
    if (s == "float" || s == "float32") return T_F32;
    if (s == "double" || s == "float64") return T_F64;
    return T_BAD;
}

struct Property {
    std::string name;
    bool list = false;
    Type count = T_BAD, item = T_BAD;
};
struct Element {
    std::string name;
    uint64_t n = 0;
    std::vector<Property> props;
};

struct Reader {
    FILE *f = nullptr;
    int format = 0;  // 0 ascii, 1 little endian, 2 big endian
    std::string err;
    ~Reader() {
        if (f) fclose(f);
    }
    bool Fail(const std::string &m) {
        if (err.empty()) err = m;
        return false;
    }
    // one value of type t as a double (rply hands every value over as a double too)
    bool Value(Type t, double *out) {
        if (format == 0) {
            char tok[64];
            if (fscanf(f, "%63s", tok) != 1) return Fail("unexpected end of file");
            char *end = nullptr;
            errno = 0;
            *out = strtod(tok, &end);
            if (end == tok || errno == ERANGE) return Fail(std::string("bad number '") + tok + "'");
            return true;
        }
        unsigned char b[8];
        const int n = kSize[t];
        if (fread(b, 1, n, f) != size_t(n)) return Fail("unexpected end of file");
        if (format == 2)
            for (int i = 0; i < n / 2; ++i) std::swap(b[i], b[n - 1 - i]);
        switch (t) {
        case T_I8: { int8_t v; memcpy(&v, b, 1); *out = v; break; }
        case T_U8: { uint8_t v; memcpy(&v, b, 1); *out = v; break; }
        case T_I16: { int16_t v; memcpy(&v, b, 2); *out = v; break; }
        case T_U16: { uint16_t v; memcpy(&v, b, 2); *out = v; break; }
        case T_I32: { int32_t v; memcpy(&v, b, 4); *out = v; break; }
        case T_U32: { uint32_t v; memcpy(&v, b, 4); *out = v; break; }
        case T_F32: { float v; memcpy(&v, b, 4); *out = v; break; }
        case T_F64: { double v; memcpy(&v, b, 8); *out = v; break; }
        default: return Fail("bad property type");
        }
        return true;
    }
};

bool ReadHeader(Reader &r, std::vector<Element> &elements) {
    char line[1024];
    auto next = [&]() -> bool { return fgets(line, sizeof line, r.f) != nullptr; };
    if (!next() || strncmp(line, "ply", 3) != 0) {
        if ((unsigned char)line[0] == 0x1f && (unsigned char)line[1] == 0x8b)
            return r.Fail("gzip-compressed PLY files are not read here: inflate the file first");
        return r.Fail("not a PLY file");
    }
    bool haveFormat = false;
    while (next()) {
        char a[64] = "", b[64] = "", c[64] = "", d[64] = "", e[64] = "";
        const int n = sscanf(line, "%63s %63s %63s %63s %63s", a, b, c, d, e);
        if (n < 1) continue;
        const std::string kw = a;
        if (kw == "end_header") return haveFormat ? true : r.Fail("PLY header without a format line");
        if (kw == "comment" || kw == "obj_info") continue;
        if (kw == "format") {
            const std::string fmt = b;
            r.format = fmt == "ascii" ? 0 : fmt == "binary_little_endian" ? 1 : fmt == "binary_big_endian" ? 2 : -1;
            if (r.format < 0) return r.Fail("unknown PLY format '" + fmt + "'");
            haveFormat = true;
        } else if (kw == "element") {
            if (n < 3) return r.Fail("bad element line in PLY header");
            Element el;
            el.name = b;
            el.n = strtoull(c, nullptr, 10);
            elements.push_back(el);
        } else if (kw == "property") {
            if (elements.empty()) return r.Fail("property before any element in PLY header");
            Property p;
            if (std::string(b) == "list") {
                if (n < 5) return r.Fail("bad list property line in PLY header");
                p.list = true;
                p.count = TypeOf(c);
                p.item = TypeOf(d);
                p.name = e;
                if (p.count == T_BAD || p.count == T_F32 || p.count == T_F64) return r.Fail("bad list count type in PLY header");
            } else {
                if (n < 3) return r.Fail("bad property line in PLY header");
                p.item = TypeOf(b);
                p.name = c;
            }
            if (p.item == T_BAD) return r.Fail("unknown property type in PLY header");
            elements.back().props.push_back(p);
        } else
            return r.Fail("unknown keyword '" + kw + "' in PLY header");
    }
    return r.Fail("PLY header without end_header");
}

template <class T>
T *Copy(const std::vector<T> &v) {
    if (v.empty()) return nullptr;
    T *p = static_cast<T *>(malloc(v.size() * sizeof(T)));
    if (p) memcpy(p, v.data(), v.size() * sizeof(T));
    return p;
}

}  // namespace

extern "C" void b200pt_ply_free(b200pt_ply_mesh *m) {
    if (!m) return;
    free(m->p);
    free(m->n);
    free(m->uv);
    free(m->tri_indices);
    free(m->quad_indices);
    free(m->face_indices);
    memset(m, 0, sizeof *m);
}

extern "C" int b200pt_ply_read(const char *path, b200pt_ply_mesh *out, char *err, uint64_t err_len) {
    auto fail = [&](const std::string &m) {
        if (err && err_len) snprintf(err, size_t(err_len), "%s: %s", path ? path : "(null)", m.c_str());
        return B200PT_ERR_INVALID;
    };
    if (!path || !out) return fail("null argument");
    memset(out, 0, sizeof *out);
    Reader r;
    r.f = fopen(path, "rb");
    if (!r.f) return fail(std::string("cannot open: ") + strerror(errno));
    std::vector<Element> elements;
    if (!ReadHeader(r, elements)) return fail(r.err);

    std::vector<float> P, N, UV;
    std::vector<int32_t> tri, quad, faceIdx;
    bool sawVertex = false, sawFace = false;
    for (const Element &el : elements) {
        const bool isVertex = el.name == "vertex", isFace = el.name == "face";
        // where each property of this element goes: >= 0 an offset into the vertex record below, -1 skipped
        enum { X, Y, Z, NX, NY, NZ, U, V, kSlots };
        std::vector<int> slot(el.props.size(), -1);
        int indicesProp = -1, faceIndexProp = -1;
        bool have[kSlots] = {};
        if (isVertex) {
            // (u, v) under the first of the reference's four conventions that the file has both halves of
            const char *uvNames[4][2] = {{"u", "v"}, {"s", "t"}, {"texture_u", "texture_v"}, {"texture_s", "texture_t"}};
            int uvPick = -1;
            for (int c = 0; c < 4 && uvPick < 0; ++c) {
                bool a = false, b = false;
                for (const Property &p : el.props) {
                    a |= !p.list && p.name == uvNames[c][0];
                    b |= !p.list && p.name == uvNames[c][1];
                }
                if (a && b) uvPick = c;
            }
            for (size_t i = 0; i < el.props.size(); ++i) {
                const Property &p = el.props[i];
                if (p.list) continue;
                const char *names[6] = {"x", "y", "z", "nx", "ny", "nz"};
                for (int k = 0; k < 6; ++k)
                    if (p.name == names[k]) slot[i] = k;
                if (uvPick >= 0 && p.name == uvNames[uvPick][0]) slot[i] = U;
                if (uvPick >= 0 && p.name == uvNames[uvPick][1]) slot[i] = V;
                if (slot[i] >= 0) have[slot[i]] = true;
            }
            if (!(have[X] && have[Y] && have[Z])) return fail("Vertex coordinate property not found!");
            sawVertex = true;
            if (el.n > (1ull << 31) - 1) return fail("too many vertices");
            P.assign(3 * el.n, 0.f);
            if (have[NX] && have[NY] && have[NZ]) N.assign(3 * el.n, 0.f);
            if (have[U] && have[V]) UV.assign(2 * el.n, 0.f);
        }
        if (isFace) {
            for (size_t i = 0; i < el.props.size(); ++i) {
                if (el.props[i].list && el.props[i].name == "vertex_indices") indicesProp = int(i);
                if (!el.props[i].list && el.props[i].name == "face_indices") faceIndexProp = int(i);
            }
            if (indicesProp < 0) return fail("vertex indices not found in PLY file");
            sawFace = true;
        }
        for (uint64_t e = 0; e < el.n; ++e) {
            for (size_t i = 0; i < el.props.size(); ++i) {
                const Property &p = el.props[i];
                double v;
                if (!p.list) {
                    if (!r.Value(p.item, &v)) return fail(r.err);
                    if (isVertex && slot[i] >= 0) {
                        const int s = slot[i];
                        if (s <= Z) P[3 * e + s] = float(v);
                        else if (s <= NZ) { if (!N.empty()) N[3 * e + (s - NX)] = float(v); }
                        else if (!UV.empty()) UV[2 * e + (s - U)] = float(v);
                    }
                    if (isFace && int(i) == faceIndexProp) faceIdx.push_back(int32_t(v));
                    continue;
                }
                if (!r.Value(p.count, &v)) return fail(r.err);
                const long len = long(v);
                if (len < 0 || len > (1 << 24)) return fail("bad list length");
                int32_t face[4] = {0, 0, 0, 0};
                for (long k = 0; k < len; ++k) {
                    if (!r.Value(p.item, &v)) return fail(r.err);
                    if (k < 4) face[k] = int32_t(v);
                }
                if (isFace && int(i) == indicesProp) {
                    if (len == 3)
                        tri.insert(tri.end(), face, face + 3);
                    else if (len == 4)  // as a bilinear patch: p00 p10 p01 p11 (util/mesh.cpp:299-305)
                        quad.insert(quad.end(), {face[0], face[1], face[3], face[2]});
                    // (other lengths: "Ignoring face with %i vertices", mesh.cpp:281-285)
                }
            }
        }
    }
    if (!sawVertex || !sawFace || P.empty() || (tri.empty() && quad.empty()))
        return fail("PLY file is invalid! No face/vertex elements found!");
    const int64_t nv = int64_t(P.size() / 3);
    for (int32_t i : tri)
        if (i < 0 || i >= nv) return fail("Vertex index " + std::to_string(i) + " is out of bounds! Valid range is [0.." + std::to_string(nv) + ")");
    for (int32_t i : quad)
        if (i < 0 || i >= nv) return fail("Vertex index " + std::to_string(i) + " is out of bounds! Valid range is [0.." + std::to_string(nv) + ")");
    out->n_vertices = uint64_t(nv);
    out->n_triangles = tri.size() / 3;
    out->n_quads = quad.size() / 4;
    out->n_face_indices = faceIdx.size();
    out->p = Copy(P);
    out->n = Copy(N);
    out->uv = Copy(UV);
    out->tri_indices = Copy(tri);
    out->quad_indices = Copy(quad);
    out->face_indices = Copy(faceIdx);
    if (!out->p || (!tri.empty() && !out->tri_indices) || (!quad.empty() && !out->quad_indices) || (!N.empty() && !out->n) ||
        (!UV.empty() && !out->uv) || (!faceIdx.empty() && !out->face_indices)) {
        b200pt_ply_free(out);
        return fail("out of memory");
    }
    return B200PT_OK;
}
