/* oracle/shim: hand-filled equivalent of the reference's
 * src/ext/openexr/OpenEXR/config/OpenEXRConfig.h.in_cmake (cmake is not run). */
#ifndef INCLUDED_OPENEXR_CONFIG_H
#define INCLUDED_OPENEXR_CONFIG_H 1
#define OPENEXR_IMF_INTERNAL_NAMESPACE_CUSTOM 0
#define OPENEXR_IMF_INTERNAL_NAMESPACE Imf_2_5
#define OPENEXR_IMF_NAMESPACE_CUSTOM 0
#define OPENEXR_IMF_NAMESPACE Imf
#define OPENEXR_VERSION_STRING "2.5.3"
#define OPENEXR_PACKAGE_STRING "OpenEXR 2.5.3"
#define OPENEXR_VERSION_MAJOR 2
#define OPENEXR_VERSION_MINOR 5
#define OPENEXR_VERSION_PATCH 3
#define OPENEXR_VERSION_HEX ((uint32_t(OPENEXR_VERSION_MAJOR) << 24) | (uint32_t(OPENEXR_VERSION_MINOR) << 16) | (uint32_t(OPENEXR_VERSION_PATCH) << 8))
#endif
