// render_host.h - b200pt_render_desc -> the POD tables of render_types.h (host copies; the API
// uploads them, tests/native/rendercheck.cpp uses them in place).  Pure host code.
#pragma once
#include <string>
#include <utility>
#include <vector>

#include "../../include/b200pt.h"
#include "render_types.h"
#include "scene_host.h"

namespace b200pt {

struct RenderHost {
    RenderSceneDev rs{};  // pointer members are filled by whoever owns the storage (host or device)
    std::vector<SpectrumTable> spectra;
    std::vector<MaterialDev> materials;
    std::vector<LightDev> lights;
    std::vector<MeshAttrDev> meshes;
    std::vector<uint32_t> lightBvhNodes;   // kLightNodeWords per node: the expanded nodes (render_lightbvh.h)
    std::vector<uint32_t> lightBitTrails;  // per light
    std::vector<int32_t> infiniteLights;   // light indices: the lights without bounds (BVHLightSampler's infiniteLights)
    std::vector<float> envImage, envDist;  // image infinite light: per-pixel spectrum coefficients, sampling distribution
    std::vector<float> filterTable;        // FilmDev::filterTable (empty: box filter)
    std::vector<float> lightGeom;          // delta lights' placement, kLightGeomFloats per light; empty = none
    std::vector<MediumDev> media;          // density / majorants pointers are filled by the owner of the grids
    std::vector<std::pair<std::vector<float>, std::vector<float>>> mediumGrids;  // per medium: density, majorants
    std::vector<std::vector<float>> mediumTempGrids;  // per medium: the temperature grid (else empty)
    std::vector<std::vector<float>> mediumLeGrids;  // per medium: the LeScale grid of an emissive grid medium (else empty)
    std::vector<float> cornerAttr;         // per-corner normals / uv (MeshAttrDev::nOff, uvOff)
    std::vector<float> lens, pupilBounds;  // realistic camera: 4 floats per element / per film-radius bin
    std::vector<int32_t> materialIds;      // GBufferFilm: the host's Material::materialId per material
    int width = 0, height = 0;
};

int BuildRenderTables(const b200pt_render_desc *d, const HostScene &hs, uint32_t nSceneMeshes, RenderHost &out,
                      std::string &err);

}  // namespace b200pt
