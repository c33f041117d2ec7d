// b200rt_resolve_noalloc: the K5 resolve (resolve.cuh arithmetic, same bytes as b200rt_resolve) for callers that
// must not allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  It has
// NO counterpart in the reference.
//
// b200rt_resolve stages the RGB8 frame in a cudaMalloc'ed buffer of the context (ctx->rgb8) between the kernel and
// the device-to-host copy.  Here that scratch is a file-scope static __device__ array of a fixed, modest size
// (B200RT_NOALLOC_MAX_PIXELS * 3 bytes = 6 MiB, held by every GPU that loads the library until exit), and the entry
// point needs no context at all: no cudaMalloc, cudaFree, stream / event creation or allocating library routine.
//
// The array is shared by every call on a device and keeps the last call's contents.  Nothing has to start at zero:
// a call overwrites bytes [0, width * height * 3) before it reads them back and never reads beyond.  Calls that
// overlap on the same GPU (two host threads, or two streams) RACE on the array: serialise them.
#include <cuda_runtime.h>
#include <stdint.h>
#include "b200rt.h"
#include "resolve.cuh"

using namespace b200rt;

// In HOST code this name is a host-side stand-in, not a device address: the kernel is given the address that
// cudaGetSymbolAddress returns (scratch_of below), never `g_noalloc_rgb` itself.
static __device__ __align__(256) uint8_t g_noalloc_rgb[(size_t)B200RT_NOALLOC_MAX_PIXELS * 3];

namespace {

#define B200RT_NOALLOC_MAX_DEVICES 64
uint8_t *g_scratch_ptr[B200RT_NOALLOC_MAX_DEVICES];     // per device, resolved once with that device current

// Device address of the scratch array on `device` (which must be the current device).
cudaError_t scratch_of(int device, uint8_t **out) {
    if (!g_scratch_ptr[device]) {
        uint8_t *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_rgb);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

__global__ void __launch_bounds__(256) resolve_noalloc_kernel(const float *__restrict__ sum, uint8_t *__restrict__ rgb, size_t n_channels,
                                                               float inv_spp, int tonemap) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_channels; i += stride)
        rgb[i] = resolve_channel(sum[i], inv_spp, tonemap);
}

}  // namespace

extern "C" int b200rt_resolve_noalloc(int32_t device, void *stream, const float *accum_sum_device, int32_t width, int32_t height,
                                      int32_t total_spp, int32_t mode, uint8_t *rgb_out) {
    if (!accum_sum_device || !rgb_out || width <= 0 || height <= 0 || total_spp <= 0) return B200RT_ERR_INVALID;
    if (mode < B200RT_MODE_HW1 || mode > B200RT_MODE_HW8) return B200RT_ERR_INVALID;
    if (device < 0 || device >= B200RT_NOALLOC_MAX_DEVICES) return B200RT_ERR_INVALID;
    if ((long long)width * height > (long long)B200RT_NOALLOC_MAX_PIXELS) return B200RT_ERR_TOO_LARGE;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) { cudaGetLastError(); return B200RT_ERR_NO_DEVICE; }
    if (device >= n_dev) return B200RT_ERR_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return B200RT_ERR_CUDA; }
    uint8_t *scratch = nullptr;
    if (scratch_of(device, &scratch) != cudaSuccess || !scratch) { cudaGetLastError(); return B200RT_ERR_CUDA; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t n = (size_t)width * height * 3;
    // same scaling and tonemap rules as b200rt_resolve: hw1 / hw2 hold one sample, hw1 is not tonemapped
    const float inv = mode <= B200RT_MODE_HW2 ? 1.f : (float)(1.0 / total_spp);
    const unsigned blocks = (unsigned)((n + 255) / 256 < 148u * 8u ? (n + 255) / 256 : 148u * 8u);
    resolve_noalloc_kernel<<<blocks, 256, 0, st>>>(accum_sum_device, scratch, n, inv, mode == B200RT_MODE_HW1 ? 0 : 1);
    if (cudaGetLastError() != cudaSuccess) return B200RT_ERR_CUDA;
    if (cudaMemcpyAsync(rgb_out, scratch, n, cudaMemcpyDeviceToHost, st) != cudaSuccess) { cudaGetLastError(); return B200RT_ERR_CUDA; }
    if (cudaStreamSynchronize(st) != cudaSuccess) { cudaGetLastError(); return B200RT_ERR_CUDA; }
    return B200RT_OK;
}
