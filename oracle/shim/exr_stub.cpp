// oracle/shim: the vendored OpenEXR library is not built for the oracle (its
// dwaLookups.h is a missing blob); every Imf symbol is left
// unresolved except this one, which InitPBRT() calls unconditionally
// (/root/reference/src/pbrt/pbrt.cpp:64).  The oracle writes PFM / raw dumps.
// TEST INFRASTRUCTURE ONLY.
#include <ImfThreading.h>
namespace OPENEXR_IMF_INTERNAL_NAMESPACE {
void setGlobalThreadCount(int) {}
}
