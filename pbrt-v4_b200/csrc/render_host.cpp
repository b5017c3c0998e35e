// render_host.cpp - validation + flattening of the render description (see render_host.h).
#include "render_host.h"
#include "render_lightbvh.h"

#include <cmath>
#include <cstring>
#include <map>

namespace b200pt {

// PiecewiseConstant2D(f, domain) (util/sampling.h:731-747) out of PiecewiseConstant1D's (:625-649), in float as there.
// Layout: f[ny][nx] | conditional cdf[ny][nx + 1] | conditional funcInt[ny] | marginal cdf[ny + 1] | marginal funcInt
static void BuildPiecewise2D(const float *f, int nx, int ny, float lox, float hix, float loy, float hiy, std::vector<float> &t) {
    t.assign(size_t(nx) * ny + size_t(ny) * (nx + 1) + ny + (ny + 1) + 1, 0.f);
    std::copy(f, f + size_t(nx) * ny, t.begin());
    float *condCdf = t.data() + size_t(nx) * ny, *condInt = condCdf + size_t(ny) * (nx + 1), *margCdf = condInt + ny;
    auto build = [](const float *func, int n, float lo, float hi, float *cdf) {  // returns funcInt
        cdf[0] = 0;
        for (int i = 1; i < n + 1; ++i) cdf[i] = cdf[i - 1] + std::abs(func[i - 1]) * (hi - lo) / n;
        float funcInt = cdf[n];
        for (int i = 1; i < n + 1; ++i) cdf[i] = funcInt == 0 ? float(i) / float(n) : cdf[i] / funcInt;
        return funcInt;
    };
    for (int v = 0; v < ny; ++v) condInt[v] = build(f + size_t(v) * nx, nx, lox, hix, condCdf + size_t(v) * (nx + 1));
    margCdf[ny + 1] = build(condInt, ny, loy, hiy, margCdf);
}

static int Log2Int(uint32_t v) {
    int l = 0;
    while (v >>= 1) ++l;
    return l;
}
static uint32_t RoundUpPow2(uint32_t v) {
    uint32_t p = 1;
    while (p < v) p <<= 1;
    return p;
}

int BuildRenderTables(const b200pt_render_desc *d, const HostScene &hs, uint32_t nSceneMeshes, RenderHost &out,
                      std::string &err) {
    auto bad = [&](const char *m) {
        err = m;
        return B200PT_ERR_INVALID;
    };
    if (!d) return bad("null render description");
    if (d->camera_type > 2) return bad("unsupported camera type (0 perspective, 1 realistic, 2 omni)");
    if (d->camera_type >= 1) {
        if (d->n_lens_elements < 1 || d->n_lens_elements > 64 || !d->lens_elements)
            return bad("realistic camera: 1..64 lens elements needed");
        if (d->n_exit_pupil_bounds < 1 || !d->exit_pupil_bounds) return bad("realistic camera: exit pupil bounds missing");
        if (!(d->film_diagonal > 0)) return bad("realistic camera: film diagonal must be positive");
        for (uint32_t i = 0; i < d->n_lens_elements; ++i)
            if (d->lens_elements[4 * i] != 0 && d->lens_elements[4 * i + 2] == 0)
                return bad("realistic camera: a curved element with eta 0");
        if (d->camera_type == 2 && d->n_microlens_elements) {
            // the reference reads elementInterfaces[i - 1].eta of the MAIN lens while tracing microlens interface i
            if (d->n_microlens_elements > d->n_lens_elements) return bad("omni camera: more microlens interfaces than lens interfaces");
            if (!d->microlens_elements || !d->microlens_offsets || d->microlens_dims[0] < 1 || d->microlens_dims[1] < 1)
                return bad("omni camera: microlens tables missing");
            if (d->microlens_simulation_radius < 0 || d->microlens_simulation_radius > 8)
                return bad("omni camera: microlens simulation radius must be 0..8");
        }
    }
    if (d->filter_type > B200PT_FILTER_TRIANGLE)
        return bad("unsupported pixel filter: box, triangle, and gaussian / mitchell / sinc through filter_table, are implemented");
    if (d->filter_type != B200PT_FILTER_BOX && d->filter_type != B200PT_FILTER_TRIANGLE) {
        const int nx = d->filter_table_size[0], ny = d->filter_table_size[1];
        if (!d->filter_table || nx < 1 || ny < 1 || nx > 4096 || ny > 4096)
            return bad("filter_type needs filter_table (the tabulated filter function) and its size");
        if (!(d->filter_radius[0] > 0) || !(d->filter_radius[1] > 0)) return bad("filter radius must be positive");
        for (int i = 0; i < nx * ny; ++i)
            if (!std::isfinite(d->filter_table[i])) return bad("filter_table holds a non-finite value");
        BuildPiecewise2D(d->filter_table, nx, ny, -d->filter_radius[0], d->filter_radius[0], -d->filter_radius[1],
                         d->filter_radius[1], out.filterTable);
    }
    if (d->film_type > 2) return bad("unsupported film type");
    if (d->film_type == 2 && (d->n_buckets != 31 || d->lambda_min != 395.f || d->lambda_max != 705.f))
        return bad("gbuffer film: 31 buckets over 395..705 nm (hard-coded in the reference, film.h:383-389)");
    if (d->film_type == 1) {
        if (d->n_buckets < 1 || d->n_buckets > 64) return bad("spectral film: n_buckets must be 1..64");
        if (d->lambda_min != 395.f || d->lambda_max != 705.f)
            return bad("spectral film: lambda_min / lambda_max must be 395 / 705 nm (the build's sampled range)");
    }
    if (d->sampler_type > B200PT_SAMPLER_INDEPENDENT) return bad("unsupported sampler type (zsobol and independent are implemented)");
    if (d->light_sampler > 2) return bad("unsupported light sampler (uniform, power and bvh are implemented)");
    if (d->light_sampler == 2 && d->n_lights && (!d->light_bit_trails || (d->n_light_bvh_nodes && !d->light_bvh_nodes)))
        return bad("the bvh light sampler needs light_bvh_nodes and light_bit_trails");
    if (d->n_meshes != nSceneMeshes) return bad("mesh attribute count differs from the geometry scene");
    if ((d->n_spectra && !d->spectra) || (d->n_materials && !d->materials) || (d->n_lights && !d->lights) ||
        (d->n_meshes && !d->meshes))
        return bad("null table in render description");
    int w = d->pixel_bounds[2] - d->pixel_bounds[0], h = d->pixel_bounds[3] - d->pixel_bounds[1];
    if (w <= 0 || h <= 0) return bad("empty pixel bounds");
    if (d->samples_per_pixel < 1 || d->max_depth < 0) return bad("bad sampler / integrator parameters");
    auto specOk = [&](int32_t s) { return s >= 0 && uint32_t(s) < d->n_spectra; };
    for (int c = 0; c < 3; ++c)
        if (!specOk(d->sensor_spectrum[c])) return bad("sensor spectrum index out of range");

    RenderSceneDev &rs = out.rs;
    rs.camera.type = d->camera_type;
    rs.camera.lens = rs.camera.pupilBounds = nullptr;
    rs.camera.nLens = rs.camera.nPupilBounds = 0;
    rs.camera.nMicro = rs.camera.microDimX = rs.camera.microDimY = rs.camera.microRadius = 0;
    rs.camera.microOffsetFromSensor = 0;
    if (d->camera_type >= 1) {
        out.lens.assign(d->lens_elements, d->lens_elements + 4 * size_t(d->n_lens_elements));
        if (d->camera_type == 2) {  // the interfaces' transforms follow the table: lens[4 n + 32 i]
            static const float ident[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
            for (uint32_t i = 0; i < d->n_lens_elements; ++i)
                for (int h = 0; h < 2; ++h) {
                    const float *m = d->lens_element_transforms ? d->lens_element_transforms + 32 * size_t(i) + 16 * h : ident;
                    out.lens.insert(out.lens.end(), m, m + 16);
                }
            if (d->n_microlens_elements) {
                const uint32_t nm = d->n_microlens_elements;
                const size_t nl = size_t(d->microlens_dims[0]) * size_t(d->microlens_dims[1]);
                out.lens.insert(out.lens.end(), d->microlens_elements, d->microlens_elements + 4 * size_t(nm));
                for (uint32_t i = 0; i < nm; ++i)
                    for (int h = 0; h < 2; ++h) {
                        const float *m = d->microlens_element_transforms ? d->microlens_element_transforms + 32 * size_t(i) + 16 * h
                                                                         : ident;
                        out.lens.insert(out.lens.end(), m, m + 16);
                    }
                out.lens.insert(out.lens.end(), d->microlens_offsets, d->microlens_offsets + 2 * nl);
                rs.camera.nMicro = int32_t(nm);
                rs.camera.microDimX = d->microlens_dims[0];
                rs.camera.microDimY = d->microlens_dims[1];
                rs.camera.microRadius = d->microlens_simulation_radius;
                rs.camera.microOffsetFromSensor = d->microlens_offset_from_sensor;
            }
        }
        out.pupilBounds.assign(d->exit_pupil_bounds, d->exit_pupil_bounds + 4 * size_t(d->n_exit_pupil_bounds));
        rs.camera.nLens = int32_t(d->n_lens_elements);
        rs.camera.nPupilBounds = int32_t(d->n_exit_pupil_bounds);
        std::memcpy(rs.camera.physicalExtent, d->physical_extent, 16);
        rs.camera.filmDiagonal = d->film_diagonal;
    }
    std::memcpy(rs.camera.cameraFromRaster, d->camera_from_raster, 64);
    std::memcpy(rs.camera.renderFromCamera, d->render_from_camera, 64);
    rs.camera.lensRadius = d->lens_radius;
    rs.camera.focalDistance = d->focal_distance;
    rs.film.type = d->film_type;
    rs.film.nBuckets = d->film_type >= 1 ? int32_t(d->n_buckets) : 0;
    if (d->film_type == 2) {
        std::memcpy(rs.film.outM, d->gbuffer_output_from_render, 64);
        std::memcpy(rs.film.outMInv, d->gbuffer_output_from_render_inv, 64);
        out.materialIds.resize(d->n_materials);
        for (uint32_t i = 0; i < d->n_materials; ++i) out.materialIds[i] = d->material_ids ? d->material_ids[i] : int32_t(i + 1);
    }
    rs.film.x0 = d->pixel_bounds[0];
    rs.film.y0 = d->pixel_bounds[1];
    rs.film.x1 = d->pixel_bounds[2];
    rs.film.y1 = d->pixel_bounds[3];
    rs.film.fullResX = d->full_resolution[0];
    rs.film.fullResY = d->full_resolution[1];
    rs.film.filterType = d->filter_type;
    const bool tabulated = d->filter_type != B200PT_FILTER_BOX && d->filter_type != B200PT_FILTER_TRIANGLE;
    rs.film.filterNx = tabulated ? d->filter_table_size[0] : 0;
    rs.film.filterNy = tabulated ? d->filter_table_size[1] : 0;
    rs.film.filterTable = nullptr;  // filled by the owner of the storage
    rs.film.filterRadiusX = d->filter_radius[0];
    rs.film.filterRadiusY = d->filter_radius[1];
    rs.film.imagingRatio = d->imaging_ratio;
    rs.film.maxComponentValue = d->max_component_value;
    std::memcpy(rs.film.outputRGBFromSensorRGB, d->output_rgb_from_sensor_rgb, 36);
    for (int c = 0; c < 3; ++c) rs.film.sensorSpectrum[c] = d->sensor_spectrum[c];
    rs.film.gbufferAlbedo = (d->film_type == 2 && d->gbuffer_albedo) ? 1u : 0u;
    rs.film.albedoIlluminant = 0;
    for (int c = 0; c < 3; ++c) rs.film.xyzSpectrum[c] = 0;
    std::memset(rs.film.rgbFromXYZ, 0, sizeof(rs.film.rgbFromXYZ));
    if (rs.film.gbufferAlbedo) {
        for (int c = 0; c < 3; ++c) {
            if (!specOk(d->gbuffer_xyz_spectrum[c])) return bad("gbuffer_albedo: gbuffer_xyz_spectrum index out of range");
            rs.film.xyzSpectrum[c] = d->gbuffer_xyz_spectrum[c];
        }
        if (!specOk(d->gbuffer_illuminant)) return bad("gbuffer_albedo: gbuffer_illuminant index out of range");
        rs.film.albedoIlluminant = d->gbuffer_illuminant;
        std::memcpy(rs.film.rgbFromXYZ, d->gbuffer_rgb_from_xyz, 36);
    }
    // ZSobolSampler constructor (samplers.h:228-238)
    rs.sampler.type = d->sampler_type;  // 1: IndependentSampler, which only uses samplesPerPixel and seed
    rs.sampler.samplesPerPixel = d->samples_per_pixel;
    rs.sampler.seed = d->seed;
    rs.sampler.randomize = d->randomize;
    rs.sampler.log2Spp = Log2Int(uint32_t(d->samples_per_pixel));
    int res = int(RoundUpPow2(uint32_t(std::max(d->full_resolution[0], d->full_resolution[1]))));
    rs.sampler.nBase4Digits = Log2Int(uint32_t(res)) + (rs.sampler.log2Spp + 1) / 2;
    rs.integrator.maxDepth = d->max_depth;
    rs.integrator.regularize = d->regularize;
    rs.integrator.lightSampler = d->light_sampler;
    rs.integrator.type = d->integrator_type;
    if (d->integrator_type > 1) return bad("unsupported integrator type (0 path, 1 volpath)");
    rs.camera.medium = -1;
    rs.nMedia = 0;
    rs.media = nullptr;
    if (d->integrator_type == 1) {
        if (d->n_media && !d->media) return bad("null media table");
        if (d->camera_medium < -1 || d->camera_medium >= int32_t(d->n_media)) return bad("camera medium out of range");
        rs.camera.medium = d->camera_medium;
        rs.nMedia = d->n_media;
        out.media.resize(d->n_media);
        out.mediumGrids.resize(d->n_media);
        out.mediumLeGrids.resize(d->n_media);
        out.mediumTempGrids.resize(d->n_media);
        for (uint32_t i = 0; i < d->n_media; ++i) {
            const b200pt_medium &m = d->media[i];
            if (m.type > B200PT_MEDIUM_GRID) return bad("unsupported medium type");
            if (!specOk(m.sigma_a) || !specOk(m.sigma_s)) return bad("medium spectrum index out of range");
            MediumDev &o = out.media[i];
            std::memset(&o, 0, sizeof o);
            o.type = m.type;
            o.sigmaA = m.sigma_a;
            o.sigmaS = m.sigma_s;
            o.g = m.g;
            if (m.type == B200PT_MEDIUM_GRID) {
                if (m.nx < 1 || m.ny < 1 || m.nz < 1 || !m.density) return bad("grid medium without a density grid");
                if (m.mres[0] < 1 || m.mres[1] < 1 || m.mres[2] < 1 || !m.majorants)
                    return bad("grid medium without a majorant grid");
                std::memcpy(o.bounds, m.bounds, 24);
                std::memcpy(o.mediumFromRender, m.medium_from_render, 64);
                std::memcpy(o.renderFromMedium, m.render_from_medium, 64);
                o.nx = m.nx; o.ny = m.ny; o.nz = m.nz;
                for (int c = 0; c < 3; ++c) o.mres[c] = m.mres[c];
                out.mediumGrids[i].first.assign(m.density, m.density + size_t(m.nx) * m.ny * m.nz);
                out.mediumGrids[i].second.assign(m.majorants, m.majorants + size_t(m.mres[0]) * m.mres[1] * m.mres[2]);
            }
            if (m.emissive) {
                if (!specOk(m.le_spectrum)) return bad("emissive medium: le_spectrum index out of range");
                o.emissive = 1;
                o.le = m.le_spectrum;
                if (m.type == B200PT_MEDIUM_GRID) {
                    if (m.le_nx < 1 || m.le_ny < 1 || m.le_nz < 1 || !m.le_scale) return bad("emissive grid medium without an LeScale grid");
                    o.lnx = m.le_nx; o.lny = m.le_ny; o.lnz = m.le_nz;
                    out.mediumLeGrids[i].assign(m.le_scale, m.le_scale + size_t(m.le_nx) * m.le_ny * m.le_nz);
                    if (m.temperature) {
                        out.mediumTempGrids[i].assign(m.temperature, m.temperature + size_t(m.nx) * m.ny * m.nz);
                        o.temperatureScale = m.temperature_scale;
                        o.temperatureOffset = m.temperature_offset;
                    }
                }
            }
        }
    }

    out.spectra.resize(d->n_spectra);
    for (uint32_t i = 0; i < d->n_spectra; ++i) {
        out.spectra[i].interp = d->spectra[i].interp;
        std::memcpy(out.spectra[i].values, d->spectra[i].values, sizeof(float) * kSpecTable);
    }
    out.materials.resize(d->n_materials);
    for (uint32_t i = 0; i < d->n_materials; ++i) {
        const b200pt_material &m = d->materials[i];
        auto opt = [&](int32_t sp) { return sp < 0 || specOk(sp); };
        switch (m.type) {
        case B200PT_MAT_DIFFUSE:
        case B200PT_MAT_DIELECTRIC:
            if (!specOk(m.spectrum[0])) return bad("material spectrum index out of range");
            break;
        case B200PT_MAT_CONDUCTOR:  // eta + k, or reflectance (materials.h:498-507)
            if (!opt(m.spectrum[0]) || !opt(m.spectrum[1]) || !opt(m.spectrum[2]))
                return bad("material spectrum index out of range");
            if (m.spectrum[0] >= 0 ? m.spectrum[1] < 0 : m.spectrum[2] < 0)
                return bad("conductor needs eta and k, or reflectance");
            break;
        case B200PT_MAT_COATED_DIFFUSE: {  // reflectance, albedo (must be black: no medium between the layers), eta
            if (m.params[8] != 0) {  // coated conductor: {conductor eta | -1, k, interface eta, reflectance}; albedo is not handed over
                if (!opt(m.spectrum[0]) || !opt(m.spectrum[1]) || !specOk(m.spectrum[2]) || !opt(m.spectrum[3]))
                    return bad("material spectrum index out of range");
                if (m.spectrum[0] >= 0 ? m.spectrum[1] < 0 : m.spectrum[3] < 0)
                    return bad("coatedconductor needs conductor eta and k, or reflectance");
                if (m.params[6] < 1 || m.params[6] > 24 || m.params[7] < 1)
                    return bad("coatedconductor maxdepth must be 1..24 and nsamples >= 1");
                break;
            }
            if (!specOk(m.spectrum[0]) || !specOk(m.spectrum[2]) || !opt(m.spectrum[1]))
                return bad("material spectrum index out of range");
            if (m.spectrum[1] >= 0)
                for (int k = 0; k < B200PT_SPECTRUM_SAMPLES; ++k)
                    if (d->spectra[m.spectrum[1]].values[k] != 0)
                        return bad("coateddiffuse with a scattering medium between the layers (albedo != 0) is not implemented");
            if (m.params[6] < 1 || m.params[6] > 24 || m.params[7] < 1)
                return bad("coateddiffuse maxdepth must be 1..24 and nsamples >= 1");
            break;
        }
        default:
            return bad("unsupported material type");
        }
        out.materials[i].type = m.type;
        std::memcpy(out.materials[i].spectrum, m.spectrum, 16);
        std::memcpy(out.materials[i].params, m.params, sizeof(m.params));
    }
    // (mesh, tri) -> primitive slot of the geometry scene
    std::map<uint64_t, uint32_t> slotOf;
    for (uint32_t s = 0; s < hs.prims.size(); ++s)
        if (hs.prims[s].mesh < kInstanceTag) slotOf[(uint64_t(hs.prims[s].mesh) << 32) | hs.prims[s].tri] = s;
    // triangles per mesh, for the per-corner attribute arrays
    std::vector<uint32_t> meshTris(d->n_meshes, 0), meshKind(d->n_meshes, 0);
    for (const PrimSlot &ps : hs.prims)
        if (ps.mesh < d->n_meshes) {
            meshTris[ps.mesh] = std::max(meshTris[ps.mesh], ps.tri + 1);
            meshKind[ps.mesh] = ps.kind;
        }
    out.meshes.resize(d->n_meshes);
    for (uint32_t i = 0; i < d->n_meshes; ++i) {
        const b200pt_mesh_attr &m = d->meshes[i];
        if (m.material >= int32_t(d->n_materials)) return bad("mesh material index out of range");
        if (m.light_first >= int32_t(d->n_lights)) return bad("mesh light index out of range");
        int32_t mi = -1, mo = -1;
        if (d->integrator_type == 1 && d->mesh_media) {
            mi = d->mesh_media[2 * i];
            mo = d->mesh_media[2 * i + 1];
            if (mi < -1 || mo < -1 || mi >= int32_t(d->n_media) || mo >= int32_t(d->n_media))
                return bad("mesh medium index out of range");
        }
        uint32_t nOff = kNoCornerAttr, uvOff = kNoCornerAttr, sOff = kNoCornerAttr;
        if (m.corner_n || m.corner_uv || m.corner_s) {
            const bool patch = (meshKind[i] & 1) != 0;  // four corners per patch, normals only
            if (out.cornerAttr.size() + 29 * size_t(meshTris[i]) >= size_t(kNoCornerAttr)) return bad("too many per-corner attributes");
            if (m.corner_n) {
                nOff = uint32_t(out.cornerAttr.size());
                out.cornerAttr.insert(out.cornerAttr.end(), m.corner_n, m.corner_n + (patch ? 12 : 9) * size_t(meshTris[i]));
            }
            if (m.corner_uv) {  // three (s, t) per triangle, four per patch
                uvOff = uint32_t(out.cornerAttr.size());
                out.cornerAttr.insert(out.cornerAttr.end(), m.corner_uv, m.corner_uv + (patch ? 8 : 6) * size_t(meshTris[i]));
            }
            if (m.corner_s && !patch) {
                sOff = uint32_t(out.cornerAttr.size());
                out.cornerAttr.insert(out.cornerAttr.end(), m.corner_s, m.corner_s + 9 * size_t(meshTris[i]));
            }
        }
        out.meshes[i] = MeshAttrDev{m.material, m.light_first, m.flip_normal, nOff, mi, mo, uvOff, sOff};
    }
    out.rs.geomExt = (hs.hasPatches || !out.cornerAttr.empty()) ? 1u : 0u;
    // ... or one of the less common shading features: the material variants, the independent sampler, delta lights
    for (const MaterialDev &m : out.materials)
        if (((m.type == MAT_DIFFUSE || m.type == MAT_DIELECTRIC) && m.params[4] != 0) || (m.type == MAT_COATED_DIFFUSE && m.params[8] != 0))
            out.rs.geomExt = 1u;
    if (d->sampler_type != B200PT_SAMPLER_ZSOBOL) out.rs.geomExt = 1u;
    if (out.rs.film.gbufferAlbedo) out.rs.geomExt = 1u;  // the albedo estimate lives in the EXT = 1 shading kernels
    for (uint32_t i = 0; i < d->n_lights; ++i)
        if (d->lights[i].type >= B200PT_LIGHT_POINT) out.rs.geomExt = 1u;
    out.lights.resize(d->n_lights);
    out.rs.envLight = -1;
    out.rs.envRes = 0;
    out.rs.envImage = out.rs.envDist = nullptr;  // filled by the owner of the storage
    for (uint32_t i = 0; i < d->n_lights; ++i) {
        const b200pt_light &l = d->lights[i];
        if (l.type > B200PT_LIGHT_IMAGE_INFINITE) return bad("unsupported light type");
        if (l.type == B200PT_LIGHT_IMAGE_INFINITE) {
            if (out.rs.envLight >= 0) return bad("one image infinite light per scene");
            const int res = d->env_resolution;
            if (!d->env_image || !d->env_distribution || res < 1 || res > 16384)
                return bad("an image infinite light needs env_image, env_distribution and env_resolution");
            for (size_t k = 0; k < size_t(res) * res; ++k)
                if (!std::isfinite(d->env_distribution[k]) || d->env_distribution[k] < 0) return bad("env_distribution must be finite and >= 0");
            for (size_t k = 0; k < 4 * size_t(res) * res; ++k)
                if (std::isnan(d->env_image[k])) return bad("env_image holds a NaN");  // (c2 = -inf is what a black pixel gets)
            out.rs.envLight = int32_t(i);
            out.rs.envRes = res;
            out.envImage.assign(d->env_image, d->env_image + 4 * size_t(res) * res);
            BuildPiecewise2D(d->env_distribution, res, res, 0.f, 1.f, 0.f, 1.f, out.envDist);
        }
        if (l.type >= B200PT_LIGHT_POINT) {
            if (!d->light_geometry) return bad("point / spot / distant lights need light_geometry");
            if (out.lightGeom.empty()) out.lightGeom.assign(size_t(kLightGeomFloats) * d->n_lights, 0.f);
            const float *g = d->light_geometry + size_t(B200PT_LIGHT_GEOMETRY_FLOATS) * i;
            for (int k = 0; k < kLightGeomFloats; ++k)
                if (!std::isfinite(g[k])) return bad("light_geometry holds a non-finite value");
            std::copy(g, g + kLightGeomFloats, out.lightGeom.begin() + size_t(kLightGeomFloats) * i);
        }
        uint32_t slot = 0;
        if (l.type == B200PT_LIGHT_AREA_TRIANGLE) {
            auto it = slotOf.find((uint64_t(l.mesh) << 32) | l.tri);
            if (it == slotOf.end()) return bad("light refers to a triangle that is not in the scene");
            slot = it->second;
        } else if (l.type == B200PT_LIGHT_UNIFORM_INFINITE || l.type == B200PT_LIGHT_DISTANT || l.type == B200PT_LIGHT_IMAGE_INFINITE)
            out.infiniteLights.push_back(int32_t(i));  // Light::Bounds() is empty (lightsamplers.cpp:114-123)
        if (!specOk(l.spectrum)) return bad("light spectrum index out of range");
        if (l.alias >= int32_t(d->n_lights)) return bad("alias index out of range");
        LightDev &o = out.lights[i];
        o.type = l.type;
        o.slot = slot;
        o.spectrum = l.spectrum;
        o.scale = l.scale;
        o.twoSided = l.two_sided;
        o.cosFalloffEnd = l.cos_falloff_end;
        o.tanFalloffEnd = l.tan_falloff_end;
        o.normalizeFalloffEnd = l.normalize_falloff_end;
        o.aliasQ = l.alias_q;
        o.aliasP = l.alias_p;
        o.alias = l.alias;
        o.flipNormal = l.mesh < d->n_meshes ? d->meshes[l.mesh].flip_normal : 0;
    }
    rs.nLightBvhNodes = 0;
    if (d->light_sampler == 2) {
        const uint32_t *nodes = static_cast<const uint32_t *>(d->light_bvh_nodes);
        for (uint32_t i = 0; i < d->n_light_bvh_nodes; ++i) {
            uint32_t link = nodes[8 * i + 6];
            if (link >> 31 ? (link & 0x7FFFFFFFu) >= d->n_lights
                           : ((link & 0x7FFFFFFFu) >= d->n_light_bvh_nodes || i + 1 >= d->n_light_bvh_nodes))
                return bad("light BVH node refers outside its tables");
        }
        // the device table: what CompactLightBounds::Importance derives from a node alone, worked out here once
        // (render_lightbvh.h ExpandLightBvhNode; same float operations, this file is built without contraction)
        out.lightBvhNodes.assign(size_t(kLightNodeWords) * d->n_light_bvh_nodes, 0u);
        for (uint32_t i = 0; i < d->n_light_bvh_nodes; ++i) {
            float e[kLightNodeWords];
            ExpandLightBvhNode(nodes + 8 * size_t(i), d->light_bvh_bounds, e);
            std::memcpy(&out.lightBvhNodes[size_t(kLightNodeWords) * i], e, sizeof e);
        }
        out.lightBitTrails.assign(d->light_bit_trails, d->light_bit_trails + d->n_lights);
        rs.nLightBvhNodes = d->n_light_bvh_nodes;
        std::memcpy(rs.lightBvhBounds, d->light_bvh_bounds, 24);
    }
    rs.nInfiniteLights = uint32_t(out.infiniteLights.size());
    rs.nSpectra = d->n_spectra;
    rs.nMaterials = d->n_materials;
    rs.nLights = d->n_lights;
    rs.nMeshes = d->n_meshes;
    out.width = w;
    out.height = h;
    return B200PT_OK;
}

}  // namespace b200pt
