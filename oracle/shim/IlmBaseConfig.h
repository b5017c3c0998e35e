/* oracle/shim: hand-filled equivalent of the reference's
 * src/ext/openexr/IlmBase/config/IlmBaseConfig.h.in_cmake (cmake is not run). */
#ifndef INCLUDED_ILMBASE_CONFIG_H
#define INCLUDED_ILMBASE_CONFIG_H 1
#define HAVE_PTHREAD 1
#define HAVE_POSIX_SEMAPHORES 1
#define ILMBASE_INTERNAL_NAMESPACE_CUSTOM 0
#define IMATH_INTERNAL_NAMESPACE Imath_2_5
#define IEX_INTERNAL_NAMESPACE Iex_2_5
#define ILMTHREAD_INTERNAL_NAMESPACE IlmThread_2_5
#define ILMBASE_NAMESPACE_CUSTOM 0
#define IMATH_NAMESPACE Imath
#define IEX_NAMESPACE Iex
#define ILMTHREAD_NAMESPACE IlmThread
#define ILMBASE_VERSION_STRING "2.5.3"
#define ILMBASE_PACKAGE_STRING "IlmBase 2.5.3"
#define ILMBASE_VERSION_MAJOR 2
#define ILMBASE_VERSION_MINOR 5
#define ILMBASE_VERSION_PATCH 3
#define ILMBASE_VERSION_HEX ((uint32_t(ILMBASE_VERSION_MAJOR) << 24) | (uint32_t(ILMBASE_VERSION_MINOR) << 16) | (uint32_t(ILMBASE_VERSION_PATCH) << 8))
#endif
