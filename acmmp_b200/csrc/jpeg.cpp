// jpeg.cpp - JPEG decode on the device through nvJPEG, so that the
// command line can read the reference's images/%08d.jpg without OpenCV.
// Reference: cv::imread(path, IMREAD_GRAYSCALE) for PatchMatch (ACMMP.cpp:534-541,
// acmmp_definitions.cpp:219) - libjpeg hands back the luma plane directly for a grey read, so
// mode 0 asks nvJPEG for NVJPEG_OUTPUT_Y; cv::imread(IMREAD_COLOR) for the fusion colours
// (acmmp_definitions.cpp:857) - mode 1 = interleaved BGR.
// Decode parity (H7): nvJPEG's IDCT and chroma upsampling are not libjpeg-turbo's; the test
// bounds the luma difference against OpenCV's decode (tests/test_gpu_jpeg.py).
#include <cuda_runtime.h>
#include <nvjpeg.h>

#include <cstdint>
#include <cstdio>
#include <mutex>
#include <string>

#include "../../include/acmmp_b200.h"

namespace {
std::mutex g_mu;
nvjpegHandle_t g_handle = nullptr;
nvjpegJpegState_t g_state = nullptr;
thread_local std::string g_jpeg_error;

int jfail(int code, const std::string &msg) {
    g_jpeg_error = msg;
    return code;
}
}  // namespace

extern "C" {

const char *acmmp_jpeg_last_error(void) { return g_jpeg_error.c_str(); }

int acmmp_jpeg_info(const uint8_t *jpeg, size_t nbytes, int *width, int *height, int *components) {
    if (!jpeg || nbytes < 4) return jfail(ACMMP_ERR_INVALID, "acmmp_jpeg_info: empty input");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_handle && nvjpegCreateSimple(&g_handle) != NVJPEG_STATUS_SUCCESS) return jfail(ACMMP_ERR_CUDA, "nvjpegCreateSimple failed");
    int nc = 0, w[NVJPEG_MAX_COMPONENT] = {0}, h[NVJPEG_MAX_COMPONENT] = {0};
    nvjpegChromaSubsampling_t ss;
    if (nvjpegGetImageInfo(g_handle, jpeg, nbytes, &nc, &ss, w, h) != NVJPEG_STATUS_SUCCESS)
        return jfail(ACMMP_ERR_INVALID, "not a JPEG stream nvJPEG can parse");
    if (width) *width = w[0];
    if (height) *height = h[0];
    if (components) *components = nc;
    return 0;
}

int acmmp_decode_jpeg(int device, const uint8_t *jpeg, size_t nbytes, int mode, uint8_t *out, size_t capacity, int *width, int *height) {
    int w = 0, h = 0, nc = 0;
    int rc = acmmp_jpeg_info(jpeg, nbytes, &w, &h, &nc);
    if (rc) return rc;
    if (mode != 0 && mode != 1) return jfail(ACMMP_ERR_INVALID, "mode must be 0 (grey) or 1 (BGR)");
    const size_t ch = mode == 0 ? 1 : 3, bytes = (size_t)w * h * ch;
    if (width) *width = w;
    if (height) *height = h;
    if (!out) return 0;
    if (capacity < bytes) return jfail(ACMMP_ERR_INVALID, "output buffer too small");
    std::lock_guard<std::mutex> lock(g_mu);
    if (cudaSetDevice(device) != cudaSuccess) return jfail(ACMMP_ERR_CUDA, "cudaSetDevice failed");
    if (!g_state && nvjpegJpegStateCreate(g_handle, &g_state) != NVJPEG_STATUS_SUCCESS) return jfail(ACMMP_ERR_CUDA, "nvjpegJpegStateCreate failed");
    unsigned char *d = nullptr;
    if (cudaMalloc(reinterpret_cast<void **>(&d), bytes) != cudaSuccess) return jfail(ACMMP_ERR_CUDA, "cudaMalloc failed");
    nvjpegImage_t img;
    for (int c = 0; c < NVJPEG_MAX_COMPONENT; ++c) {
        img.channel[c] = nullptr;
        img.pitch[c] = 0;
    }
    img.channel[0] = d;
    img.pitch[0] = (size_t)w * ch;
    const nvjpegStatus_t st = nvjpegDecode(g_handle, g_state, jpeg, nbytes, mode == 0 ? NVJPEG_OUTPUT_Y : NVJPEG_OUTPUT_BGRI, &img, nullptr);
    cudaError_t e = cudaSuccess;
    if (st == NVJPEG_STATUS_SUCCESS) {
        e = cudaStreamSynchronize(nullptr);
        if (e == cudaSuccess) e = cudaMemcpy(out, d, bytes, cudaMemcpyDeviceToHost);
    }
    cudaFree(d);
    if (st != NVJPEG_STATUS_SUCCESS) return jfail(ACMMP_ERR_CUDA, "nvjpegDecode failed with status " + std::to_string((int)st));
    if (e != cudaSuccess) return jfail(ACMMP_ERR_CUDA, std::string("copy of the decoded image failed: ") + cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
