// exr_pack.cu - film finalisation straight into the payload of an OpenEXR file (the reference packs on
// the CPU: Film::WriteImage -> Image::WriteEXR, film.cpp:542-574, util/image.cpp:1008+).  The image kernels leave a
// channels-last float image on the device; this kernel turns it into the scanline chunks of an uncompressed scanline
// EXR - per line {int32 y, int32 bytes} followed by the line of every channel in the file's (alphabetical) channel
// order, half or float - so that the host only prepends the header and the offset table and writes one buffer.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>

namespace b200pt {

// order[c] = index into the channels-last image of the c-th channel of the file
__global__ void ExrPackKernel(const float *__restrict__ img, int width, int height, int nChannels, const int *__restrict__ order,
                              int half, int y0, unsigned char *__restrict__ out) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long perLine = (long long)nChannels * width;
    if (i >= perLine * height) return;
    const int y = (int)(i / perLine), c = (int)((i % perLine) / width), x = (int)(i % width);
    const int bytes = half ? 2 : 4;
    const long long lineBytes = perLine * bytes;
    unsigned char *line = out + (long long)y * (8 + lineBytes);
    if (c == 0 && x == 0) {
        reinterpret_cast<int *>(line)[0] = y0 + y;
        reinterpret_cast<int *>(line)[1] = (int)lineBytes;
    }
    float v = img[((long long)y * width + x) * nChannels + order[c]];
    unsigned char *dst = line + 8 + ((long long)c * width + x) * bytes;
    if (half) {
        if (v > 65504.f) v = 65504.f;  // as the reference clamps before writing fp16 (film.cpp:871-879)
        *reinterpret_cast<__half *>(dst) = __float2half_rn(v);
    } else {
        *reinterpret_cast<float *>(dst) = v;
    }
}

cudaError_t LaunchExrPack(const float *img, int width, int height, int nChannels, const int *order, bool half, int y0,
                          unsigned char *out, cudaStream_t stream) {
    const long long n = (long long)width * height * nChannels;
    ExrPackKernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(img, width, height, nChannels, order, half ? 1 : 0, y0, out);
    return cudaGetLastError();
}

}  // namespace b200pt
