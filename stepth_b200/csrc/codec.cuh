// codec.cuh -- SURVEY 8(f) row 4: decode / encode on the device, for hosts that want the images to arrive in HBM already.
//
// The reference decodes with `image::open` (src/depth_image.rs:14,66,81; jpeg-decoder 0.1.22 underneath) and saves with
// `save` (src/mask_image.rs:197-199, DepthImage::depth().save()).  This is NOT on the parity path: a JPEG decoder's IDCT and
// chroma upsampling are implementation-defined to within a level or two, so pixels decoded here differ from jpeg-decoder's --
// a drop-in that must match the reference bit for bit keeps decoding on the host, as rust/stepth does.  What this offers is
// throughput: nvJPEG (a CUDA toolkit library, resolved at run time like NCCL -- a box without it still loads the library and
// every other entry point) decodes a baseline / progressive JPEG into interleaved RGB8 in device memory and encodes RGB8 /
// RGBA8 / Luma8 planes.  Library code, called, not written here: it counts like a cuBLAS call.
#pragma once
#include <dlfcn.h>
#include <nvjpeg.h>

#include <mutex>

#include "ctx.cuh"


namespace stepth {

struct NvjpegApi {
    void *lib = nullptr;
    decltype(&nvjpegCreateSimple) CreateSimple = nullptr;
    decltype(&nvjpegDestroy) Destroy = nullptr;
    decltype(&nvjpegJpegStateCreate) JpegStateCreate = nullptr;
    decltype(&nvjpegJpegStateDestroy) JpegStateDestroy = nullptr;
    decltype(&nvjpegGetImageInfo) GetImageInfo = nullptr;
    decltype(&nvjpegDecode) Decode = nullptr;
    decltype(&nvjpegEncoderStateCreate) EncoderStateCreate = nullptr;
    decltype(&nvjpegEncoderStateDestroy) EncoderStateDestroy = nullptr;
    decltype(&nvjpegEncoderParamsCreate) EncoderParamsCreate = nullptr;
    decltype(&nvjpegEncoderParamsDestroy) EncoderParamsDestroy = nullptr;
    decltype(&nvjpegEncoderParamsSetQuality) EncoderParamsSetQuality = nullptr;
    decltype(&nvjpegEncoderParamsSetSamplingFactors) EncoderParamsSetSamplingFactors = nullptr;
    decltype(&nvjpegEncodeImage) EncodeImage = nullptr;
    decltype(&nvjpegEncodeYUV) EncodeYUV = nullptr;
    decltype(&nvjpegEncodeRetrieveBitstream) EncodeRetrieveBitstream = nullptr;
    std::string why;
};

inline NvjpegApi &nvjpeg_api() {
    static NvjpegApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char *name : {"libnvjpeg.so.12", "libnvjpeg.so", "/usr/local/cuda/lib64/libnvjpeg.so.12"}) {
            api.lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
            if (api.lib) break;
        }
        if (!api.lib) {
            api.why = std::string("nvJPEG not found (dlopen libnvjpeg.so.12): ") + (dlerror() ? dlerror() : "?");
            return;
        }
        bool ok = true;
        auto sym = [&](const char *n) {
            void *p = dlsym(api.lib, n);
            if (!p) {
                ok = false;
                api.why = std::string("nvJPEG symbol missing: ") + n;
            }
            return p;
        };
#define STEPTH_NVJ(field, name) api.field = (decltype(api.field))sym(name)
        STEPTH_NVJ(CreateSimple, "nvjpegCreateSimple");
        STEPTH_NVJ(Destroy, "nvjpegDestroy");
        STEPTH_NVJ(JpegStateCreate, "nvjpegJpegStateCreate");
        STEPTH_NVJ(JpegStateDestroy, "nvjpegJpegStateDestroy");
        STEPTH_NVJ(GetImageInfo, "nvjpegGetImageInfo");
        STEPTH_NVJ(Decode, "nvjpegDecode");
        STEPTH_NVJ(EncoderStateCreate, "nvjpegEncoderStateCreate");
        STEPTH_NVJ(EncoderStateDestroy, "nvjpegEncoderStateDestroy");
        STEPTH_NVJ(EncoderParamsCreate, "nvjpegEncoderParamsCreate");
        STEPTH_NVJ(EncoderParamsDestroy, "nvjpegEncoderParamsDestroy");
        STEPTH_NVJ(EncoderParamsSetQuality, "nvjpegEncoderParamsSetQuality");
        STEPTH_NVJ(EncoderParamsSetSamplingFactors, "nvjpegEncoderParamsSetSamplingFactors");
        STEPTH_NVJ(EncodeImage, "nvjpegEncodeImage");
        STEPTH_NVJ(EncodeYUV, "nvjpegEncodeYUV");
        STEPTH_NVJ(EncodeRetrieveBitstream, "nvjpegEncodeRetrieveBitstream");
#undef STEPTH_NVJ
        if (!ok) {
            dlclose(api.lib);
            api.lib = nullptr;
        }
    });
    return api;
}

#define STEPTH_NVJ_TRY(ctx, expr)                                                                                           \
    do {                                                                                                                    \
        nvjpegStatus_t r__ = (expr);                                                                                        \
        if (r__ != NVJPEG_STATUS_SUCCESS)                                                                                   \
            return stepth::set_err((ctx), STEPTH_ERR_CODEC, "%s failed: nvjpeg status %d (%s:%d)", #expr, (int)r__, __FILE__, \
                                   __LINE__);                                                                               \
    } while (0)

// the codec objects of a context, created on first use
inline int32_t codec_ensure(stepth_ctx *ctx) {
    NvjpegApi &nj = nvjpeg_api();
    if (!nj.lib) return set_err(ctx, STEPTH_ERR_CODEC, "%s", nj.why.c_str());
    if (ctx->nvj_handle) return STEPTH_OK;
    nvjpegHandle_t h = nullptr;
    STEPTH_NVJ_TRY(ctx, nj.CreateSimple(&h));
    ctx->nvj_handle = h;
    nvjpegJpegState_t st = nullptr;
    STEPTH_NVJ_TRY(ctx, nj.JpegStateCreate(h, &st));
    ctx->nvj_state = st;
    nvjpegEncoderState_t es = nullptr;
    STEPTH_NVJ_TRY(ctx, nj.EncoderStateCreate(h, &es, ctx->stream));
    ctx->nvj_enc_state = es;
    nvjpegEncoderParams_t ep = nullptr;
    STEPTH_NVJ_TRY(ctx, nj.EncoderParamsCreate(h, &ep, ctx->stream));
    ctx->nvj_enc_params = ep;
    return STEPTH_OK;
}

inline void codec_destroy(stepth_ctx *ctx) {
    NvjpegApi &nj = nvjpeg_api();
    if (!nj.lib) return;
    if (ctx->nvj_enc_params) nj.EncoderParamsDestroy((nvjpegEncoderParams_t)ctx->nvj_enc_params);
    if (ctx->nvj_enc_state) nj.EncoderStateDestroy((nvjpegEncoderState_t)ctx->nvj_enc_state);
    if (ctx->nvj_state) nj.JpegStateDestroy((nvjpegJpegState_t)ctx->nvj_state);
    if (ctx->nvj_handle) nj.Destroy((nvjpegHandle_t)ctx->nvj_handle);
    ctx->nvj_enc_params = ctx->nvj_enc_state = ctx->nvj_state = ctx->nvj_handle = nullptr;
}

#if defined(__CUDACC__)
// RGBA8 -> RGB8 (the encoder takes three channels; `save` of an RGBA image drops alpha the same way when the format has none)
__global__ void __launch_bounds__(256) rgba_to_rgb_kernel(const uint32_t *rgba, uint8_t *rgb, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t p = rgba[i];
        rgb[3 * i] = (uint8_t)p;
        rgb[3 * i + 1] = (uint8_t)(p >> 8);
        rgb[3 * i + 2] = (uint8_t)(p >> 16);
    }
}
#endif

}  // namespace stepth
