// nvdec.cu -- video ingest on the device: compressed packets in, frames in HBM out.
//
// The reference reads its video with cv2.VideoCapture(path).read() (src/tests/tester_map.py:58-66, run_video.py:
// 138-149): FFmpeg decodes on the CPU, swscale converts YUV 4:2:0 to BGR24, and the extractor converts BGR to gray
// (extractor.py:61-62) -- three host passes and a PCIe upload per frame in front of a front end that runs at > 1e5
// frames/s.  Here the compressed packets (kilobytes) go to the GPU's NVDEC engine and the decoded NV12 surface is
// turned into the extractor's input by one kernel, without a host hop:
//     packets -> cuvid parser -> NVDEC -> NV12 surface -> vs_nv12_kernel -> gray (or BGR) frame in HBM.
// The demuxer stays on the host (container parsing is IO, not compute): cv2.VideoCapture in raw mode
// (CAP_PROP_FORMAT = -1) hands out the encoded packets, see orbslam2/ingest.py.
//
// libnvcuvid ships with the driver, not with the CUDA toolkit, and its headers (Video Codec SDK) are not in this
// image: the library is dlopen'ed and the handful of entry points and structures used here are declared by hand,
// following the public nvcuvid.h / cuviddec.h interface (CUDA Video Decoder API).  Nothing links against it: on a box
// without the library vslam_nvdec_create fails with VSLAM_E_CUDA and a clear message.
//
// Output modes (vslam_nvdec_decode):
//   VSLAM_NVDEC_LUMA  the Y plane as is (what a native pipeline would feed the extractor; NOT what the reference
//                     computes: its gray is BGR2GRAY of swscale's BGR, i.e. range-expanded and chroma-dependent);
//   VSLAM_NVDEC_GRAY  the reference's gray, bit for bit: swscale's yuv420p -> bgr24 conversion of the decoded frame
//                     followed by OpenCV's BGR2GRAY (K0) -- see vs_swscale_bgr below for the arithmetic;
//   VSLAM_NVDEC_BGR   that BGR frame itself (uint8 [h][w][3]), what cv2.VideoCapture.read() returns.
#include <dlfcn.h>

#include <mutex>
#include <new>

#include "common.cuh"

// ---- hand-declared subset of the CUDA Video Decoder API (nvcuvid.h / cuviddec.h) -------------------------------
namespace {

typedef void* CUvideodecoder;
typedef void* CUvideoparser;
typedef void* CUvideoctxlock;
typedef long long CUvideotimestamp;
typedef int CUresult_t;

enum { cudaVideoCodec_MPEG4 = 2, cudaVideoCodec_H264 = 4, cudaVideoCodec_JPEG = 5, cudaVideoCodec_HEVC = 8,
       cudaVideoCodec_VP8 = 9, cudaVideoCodec_VP9 = 10, cudaVideoCodec_AV1 = 11 };
enum { cudaVideoChromaFormat_420 = 1 };
enum { cudaVideoSurfaceFormat_NV12 = 0 };
enum { cudaVideoDeinterlaceMode_Weave = 0 };
enum { cudaVideoCreate_PreferCUVID = 4 };
enum { CUVID_PKT_ENDOFSTREAM = 1, CUVID_PKT_ENDOFPICTURE = 8 };

struct CUVIDEOFORMAT {
    int codec;
    struct { unsigned int numerator, denominator; } frame_rate;
    unsigned char progressive_sequence, bit_depth_luma_minus8, bit_depth_chroma_minus8, min_num_decode_surfaces;
    unsigned int coded_width, coded_height;
    struct { int left, top, right, bottom; } display_area;
    int chroma_format;
    unsigned int bitrate;
    struct { int x, y; } display_aspect_ratio;
    struct { unsigned char flags, color_primaries, transfer_characteristics, matrix_coefficients; } video_signal_description;
    unsigned int seqhdr_data_length;
};

struct CUVIDPICPARAMS_HEAD {     // the parser owns the (4 KB) structure; only the head is read here
    int PicWidthInMbs, FrameHeightInMbs, CurrPicIdx;
};

struct CUVIDPARSERDISPINFO {
    int picture_index, progressive_frame, top_field_first, repeat_first_field;
    CUvideotimestamp timestamp;
};

typedef int (*PFNVIDSEQUENCECALLBACK)(void*, CUVIDEOFORMAT*);
typedef int (*PFNVIDDECODECALLBACK)(void*, void*);
typedef int (*PFNVIDDISPLAYCALLBACK)(void*, CUVIDPARSERDISPINFO*);

struct CUVIDPARSERPARAMS {
    int CodecType;
    unsigned int ulMaxNumDecodeSurfaces, ulClockRate, ulErrorThreshold, ulMaxDisplayDelay;
    unsigned int flags;                       // bAnnexb : 1, reserved : 31
    unsigned int uReserved1[4];
    void* pUserData;
    PFNVIDSEQUENCECALLBACK pfnSequenceCallback;
    PFNVIDDECODECALLBACK pfnDecodePicture;
    PFNVIDDISPLAYCALLBACK pfnDisplayPicture;
    void* pfnGetOperatingPoint;
    void* pfnGetSEIMsg;
    void* pvReserved2[5];
    void* pExtVideoInfo;
};

struct CUVIDSOURCEDATAPACKET {
    unsigned long flags, payload_size;
    const unsigned char* payload;
    CUvideotimestamp timestamp;
};

struct CUVIDDECODECREATEINFO {
    unsigned long ulWidth, ulHeight, ulNumDecodeSurfaces;
    int CodecType, ChromaFormat;
    unsigned long ulCreationFlags, bitDepthMinus8, ulIntraDecodeOnly, ulMaxWidth, ulMaxHeight, Reserved1;
    struct { short left, top, right, bottom; } display_area;
    int OutputFormat, DeinterlaceMode;
    unsigned long ulTargetWidth, ulTargetHeight, ulNumOutputSurfaces;
    CUvideoctxlock vidLock;
    struct { short left, top, right, bottom; } target_rect;
    unsigned long enableHistogram;
    unsigned long Reserved2[4];
};

struct CUVIDPROCPARAMS {
    int progressive_frame, second_field, top_field_first, unpaired_field;
    unsigned int reserved_flags, reserved_zero;
    unsigned long long raw_input_dptr;
    unsigned int raw_input_pitch, raw_input_format;
    unsigned long long raw_output_dptr;
    unsigned int raw_output_pitch, Reserved1;
    cudaStream_t output_stream;
    unsigned int Reserved[46];
    unsigned long long* histogram_dptr;
    void* Reserved2[1];
};

struct CuvidApi {
    void* lib = nullptr;
    CUresult_t (*CreateVideoParser)(CUvideoparser*, CUVIDPARSERPARAMS*) = nullptr;
    CUresult_t (*ParseVideoData)(CUvideoparser, CUVIDSOURCEDATAPACKET*) = nullptr;
    CUresult_t (*DestroyVideoParser)(CUvideoparser) = nullptr;
    CUresult_t (*CreateDecoder)(CUvideodecoder*, CUVIDDECODECREATEINFO*) = nullptr;
    CUresult_t (*DestroyDecoder)(CUvideodecoder) = nullptr;
    CUresult_t (*DecodePicture)(CUvideodecoder, void*) = nullptr;
    CUresult_t (*MapVideoFrame64)(CUvideodecoder, int, unsigned long long*, unsigned int*, CUVIDPROCPARAMS*) = nullptr;
    CUresult_t (*UnmapVideoFrame64)(CUvideodecoder, unsigned long long) = nullptr;
    CUresult_t (*GetDecoderCaps)(void*) = nullptr;
    bool ok = false;
};

CuvidApi& cuvid()
{
    static CuvidApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnvcuvid.so.1", "libnvcuvid.so"};
        for (const char* n : names)
            if ((api.lib = dlopen(n, RTLD_NOW | RTLD_LOCAL))) break;
        if (!api.lib) return;
#define VS_SYM(field, name) *(void**)&api.field = dlsym(api.lib, name)
        VS_SYM(CreateVideoParser, "cuvidCreateVideoParser");
        VS_SYM(ParseVideoData, "cuvidParseVideoData");
        VS_SYM(DestroyVideoParser, "cuvidDestroyVideoParser");
        VS_SYM(CreateDecoder, "cuvidCreateDecoder");
        VS_SYM(DestroyDecoder, "cuvidDestroyDecoder");
        VS_SYM(DecodePicture, "cuvidDecodePicture");
        VS_SYM(MapVideoFrame64, "cuvidMapVideoFrame64");
        VS_SYM(UnmapVideoFrame64, "cuvidUnmapVideoFrame64");
        VS_SYM(GetDecoderCaps, "cuvidGetDecoderCaps");
#undef VS_SYM
        api.ok = api.CreateVideoParser && api.ParseVideoData && api.DestroyVideoParser && api.CreateDecoder &&
                 api.DestroyDecoder && api.DecodePicture && api.MapVideoFrame64 && api.UnmapVideoFrame64;
    });
    return api;
}

// ---- NV12 -> the extractor's input ---------------------------------------------------------------------------
// swscale's unscaled yuv420p -> bgr24 converter (what cv2.VideoCapture's FFmpeg backend runs: same size, no
// SWS_ACCURATE_RND): per pixel, with the chroma sample of its 2x2 block (no chroma interpolation),
//     Y' = (Y << 3) - yoff,  U' = (U << 3) - 1024,  V' = (V << 3) - 1024          (16-bit lanes)
//     y  = mulhi16(Y', cy);  b = y + mulhi16(U', cbu);  g = y + mulhi16(U', cgu) + mulhi16(V', cgv);
//     r  = y + mulhi16(V', crv);  each clamped to [0, 255]                          (paddsw + packuswb)
// i.e. the SIMD path of libswscale/x86/yuv2rgb_template.c with the ITU-R BT.601 limited-range table that
// sws_getCachedContext selects by default.  The five coefficients and the luma offset are host parameters of the
// kernel (vslam_nvdec_set_colour), because they are what a different swscale build or colour space would change.
// UNVERIFIED: the defaults are the textbook 13-bit BT.601 values; they were to be fitted against cv2's BGR output on a
// box with NVDEC, and this pool's boxes do not expose the engine (profiles/r2_nvdec_probe.txt).
// tests/test_gpu_ingest.py holds the bit-equality check for the day it runs on one that does.
struct VsYuvCoef { int cy, cbu, cgu, cgv, crv, yoff; };

__device__ __forceinline__ int mulhi16(int a, int b) { return (a * b) >> 16; }     // pmulhw on sign-extended lanes
__device__ __forceinline__ int clamp255(int v) { return v < 0 ? 0 : v > 255 ? 255 : v; }

// mode 0: luma copy, 1: reference gray (swscale BGR + OpenCV BGR2GRAY), 2: BGR, 3: NV12 copy.  One thread = 4 pixels of a row.
__global__ void __launch_bounds__(256)
vs_nv12_kernel(const uint8_t* __restrict__ nv12, unsigned pitch, int luma_rows, int w, int h, VsYuvCoef k, int mode,
               uint8_t* __restrict__ out, int out_pitch)
{
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (mode == 3) {               // the decoder's own NV12: h luma rows, then h / 2 rows of interleaved U, V
        if (x >= w || y >= h + h / 2) return;
        const uint8_t* src = nv12 + (size_t)(y < h ? y : luma_rows + (y - h)) * pitch;
        for (int j = 0; j < 4 && x + j < w; ++j) out[(size_t)y * out_pitch + x + j] = src[x + j];
        return;
    }
    if (x >= w || y >= h) return;
    const uint8_t* yrow = nv12 + (size_t)y * pitch;
    const uint8_t* uvrow = nv12 + (size_t)luma_rows * pitch + (size_t)(y >> 1) * pitch;
    uint8_t g[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int xx = min(x + j, w - 1);
        const int Y = yrow[xx];
        if (mode == 0) { g[j] = (uint8_t)Y; continue; }
        const int U = uvrow[xx & ~1], V = uvrow[(xx & ~1) + 1];
        const int ys = mulhi16((Y << 3) - k.yoff, k.cy), us = (U << 3) - 1024, vs = (V << 3) - 1024;
        const int b = clamp255(ys + mulhi16(us, k.cbu));
        const int gg = clamp255(ys + mulhi16(us, k.cgu) + mulhi16(vs, k.cgv));
        const int r = clamp255(ys + mulhi16(vs, k.crv));
        if (mode == 2) {
            if (x + j < w) {
                uint8_t* o = out + (size_t)y * out_pitch + (size_t)(x + j) * 3;
                o[0] = (uint8_t)b; o[1] = (uint8_t)gg; o[2] = (uint8_t)r;
            }
        } else {
            g[j] = (uint8_t)((b * 3735 + gg * 19235 + r * 9798 + 16384) >> 15);      // K0: cv2.cvtColor(BGR2GRAY)
        }
    }
    if (mode != 2) {
        uint8_t* o = out + (size_t)y * out_pitch + x;
        if (x + 3 < w && (((size_t)o) & 3) == 0) *(uint32_t*)o = g[0] | (g[1] << 8) | (g[2] << 16) | ((uint32_t)g[3] << 24);
        else
            for (int j = 0; j < 4 && x + j < w; ++j) o[j] = g[j];
    }
}

}  // namespace

struct vslam_nvdec {
    CUvideoparser parser = nullptr;
    CUvideodecoder decoder = nullptr;
    int codec = 0, width = 0, height = 0, coded_w = 0, coded_h = 0, surfaces = 0;
    VsYuvCoef coef = {9539, 16525, -3209, -6660, 13074, 128};   // BT.601 limited range, swscale's 13-bit fixed point
    // destination of the packet being parsed (set by vslam_nvdec_decode for the duration of the call)
    uint8_t* d_out = nullptr;
    size_t frame_stride = 0;
    int out_pitch = 0, max_frames = 0, mode = 0, n_out = 0, dropped = 0;
    cudaStream_t stream = nullptr;
    int error = 0;
};

namespace {

int on_sequence(void* user, CUVIDEOFORMAT* f)
{
    vslam_nvdec* d = (vslam_nvdec*)user;
    CuvidApi& api = cuvid();
    if (f->chroma_format != cudaVideoChromaFormat_420 || f->bit_depth_luma_minus8 != 0) {
        vs_set_error("vslam_nvdec: only 8-bit 4:2:0 streams are handled (chroma format %d, bit depth %d)", f->chroma_format,
                     8 + f->bit_depth_luma_minus8);
        d->error = VSLAM_E_INVALID;
        return 0;
    }
    const int w = f->display_area.right - f->display_area.left, h = f->display_area.bottom - f->display_area.top;
    int n = f->min_num_decode_surfaces;
    if (n < 8) n = 8;
    if (d->decoder && (int)f->coded_width == d->coded_w && (int)f->coded_height == d->coded_h && w == d->width && h == d->height)
        return n;                                     // the same sequence header again
    if (d->decoder) { api.DestroyDecoder(d->decoder); d->decoder = nullptr; }
    CUVIDDECODECREATEINFO ci;
    memset(&ci, 0, sizeof(ci));
    ci.ulWidth = f->coded_width; ci.ulHeight = f->coded_height;
    ci.ulNumDecodeSurfaces = n;
    ci.CodecType = f->codec; ci.ChromaFormat = f->chroma_format;
    ci.ulCreationFlags = cudaVideoCreate_PreferCUVID;
    ci.bitDepthMinus8 = 0;
    ci.ulMaxWidth = f->coded_width; ci.ulMaxHeight = f->coded_height;
    ci.display_area.left = (short)f->display_area.left; ci.display_area.top = (short)f->display_area.top;
    ci.display_area.right = (short)f->display_area.right; ci.display_area.bottom = (short)f->display_area.bottom;
    ci.OutputFormat = cudaVideoSurfaceFormat_NV12;
    ci.DeinterlaceMode = cudaVideoDeinterlaceMode_Weave;
    ci.ulTargetWidth = w; ci.ulTargetHeight = h;
    ci.ulNumOutputSurfaces = 2;
    const CUresult_t rc = api.CreateDecoder(&d->decoder, &ci);
    if (rc != 0) {
        vs_set_error("vslam_nvdec: cuvidCreateDecoder failed with CUresult %d (codec %d, %ux%u)", rc, f->codec, f->coded_width,
                     f->coded_height);
        d->error = VSLAM_E_CUDA;
        d->decoder = nullptr;
        return 0;
    }
    d->width = w; d->height = h; d->coded_w = f->coded_width; d->coded_h = f->coded_height; d->surfaces = n;
    return n;
}

int on_decode(void* user, void* pic)
{
    vslam_nvdec* d = (vslam_nvdec*)user;
    if (!d->decoder) return 0;
    const CUresult_t rc = cuvid().DecodePicture(d->decoder, pic);
    if (rc != 0) {
        vs_set_error("vslam_nvdec: cuvidDecodePicture failed with CUresult %d", rc);
        d->error = VSLAM_E_CUDA;
        return 0;
    }
    return 1;
}

int on_display(void* user, CUVIDPARSERDISPINFO* info)
{
    vslam_nvdec* d = (vslam_nvdec*)user;
    if (!info || !d->decoder) return 1;              // end-of-stream marker
    if (d->n_out >= d->max_frames || !d->d_out) { ++d->dropped; return 1; }
    const size_t rows = d->mode == VSLAM_NVDEC_NV12 ? d->height + d->height / 2 : d->height;
    if (d->out_pitch < (d->mode == VSLAM_NVDEC_BGR ? 3 : 1) * d->width || d->frame_stride < (size_t)d->out_pitch * rows) {
        vs_set_error("vslam_nvdec_decode: output pitch / frame stride too small for %d x %d frames", d->width, d->height);
        d->error = VSLAM_E_INVALID;
        return 0;
    }
    CuvidApi& api = cuvid();
    CUVIDPROCPARAMS pp;
    memset(&pp, 0, sizeof(pp));
    pp.progressive_frame = info->progressive_frame;
    pp.top_field_first = info->top_field_first;
    pp.unpaired_field = info->repeat_first_field < 0;
    pp.output_stream = d->stream;
    unsigned long long src = 0;
    unsigned int pitch = 0;
    CUresult_t rc = api.MapVideoFrame64(d->decoder, info->picture_index, &src, &pitch, &pp);
    if (rc != 0) {
        vs_set_error("vslam_nvdec: cuvidMapVideoFrame failed with CUresult %d", rc);
        d->error = VSLAM_E_CUDA;
        return 0;
    }
    // NV12: `height` luma rows (the surface of the target size), then height / 2 rows of interleaved U, V
    const dim3 grid((d->width / 4 + 255) / 256 + 1, d->mode == VSLAM_NVDEC_NV12 ? d->height + d->height / 2 : d->height);
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_nv12_kernel<<<grid, 256, 0, d->stream>>>((const uint8_t*)src, pitch, d->height, d->width, d->height, d->coef, d->mode,
                                                 d->d_out + (size_t)d->n_out * d->frame_stride, d->out_pitch);
    const cudaError_t e = cudaGetLastError();
    // the surface returns to the decoder at unmap: the conversion has to be done with it by then
    cudaStreamSynchronize(d->stream);
    api.UnmapVideoFrame64(d->decoder, src);
    if (e != cudaSuccess) { d->error = vs_cuda_fail(e, "vs_nv12_kernel", __FILE__, __LINE__); return 0; }
    ++d->n_out;
    return 1;
}

}  // namespace

// CUVIDDECODECAPS (cuviddec.h): in = codec, chroma format, bit depth; out = bIsSupported, engine count, limits
struct CUVIDDECODECAPS {
    int eCodecType, eChromaFormat;
    unsigned int nBitDepthMinus8, reserved1[3];
    unsigned char bIsSupported, nNumNVDECs;
    unsigned short nOutputFormatMask;
    unsigned int nMaxWidth, nMaxHeight, nMaxMBCount;
    unsigned short nMinWidth, nMinHeight;
    unsigned char bIsHistogramSupported, nCounterBitDepth;
    unsigned short nMaxHistogramBins;
    unsigned int reserved3[10];
};

// 1 = the library is loaded AND the engine answers for this codec; 0 otherwise, *cu_result = what cuvid said
// (100 = CUDA_ERROR_NO_DEVICE: the container was started without the `video` driver capability, so the NVDEC
// engines are not exposed to it although libnvcuvid is on the library path -- the state of this pool's boxes)
static int nvdec_engine_ok(int cu_codec, int* cu_result)
{
    CuvidApi& api = cuvid();
    *cu_result = -1;
    if (!api.ok || !api.GetDecoderCaps) return 0;
    if (cudaFree(nullptr) != cudaSuccess) { cudaGetLastError(); return 0; }
    CUVIDDECODECAPS caps;
    memset(&caps, 0, sizeof(caps));
    caps.eCodecType = cu_codec;
    caps.eChromaFormat = cudaVideoChromaFormat_420;
    *cu_result = api.GetDecoderCaps(&caps);
    return *cu_result == 0 && caps.bIsSupported;
}

extern "C" int vslam_nvdec_available(void)
{
    int rc;
    return nvdec_engine_ok(cudaVideoCodec_H264, &rc) || nvdec_engine_ok(cudaVideoCodec_VP9, &rc);
}

extern "C" int vslam_nvdec_create(int codec, vslam_nvdec** out)
{
    VS_REQUIRE(out, "null argument");
    *out = nullptr;
    CuvidApi& api = cuvid();
    if (!api.ok) {
        vs_set_error("vslam_nvdec_create: libnvcuvid.so.1 (NVDEC driver library) is not loadable on this box: %s",
                     api.lib ? "symbols missing" : "dlopen failed");
        return VSLAM_E_CUDA;
    }
    int cu = 0;
    switch (codec) {
    case VSLAM_CODEC_MPEG4: cu = cudaVideoCodec_MPEG4; break;
    case VSLAM_CODEC_H264: cu = cudaVideoCodec_H264; break;
    case VSLAM_CODEC_MJPEG: cu = cudaVideoCodec_JPEG; break;
    case VSLAM_CODEC_HEVC: cu = cudaVideoCodec_HEVC; break;
    case VSLAM_CODEC_VP8: cu = cudaVideoCodec_VP8; break;
    case VSLAM_CODEC_VP9: cu = cudaVideoCodec_VP9; break;
    case VSLAM_CODEC_AV1: cu = cudaVideoCodec_AV1; break;
    default:
        vs_set_error("vslam_nvdec_create: unknown codec %d", codec);
        return VSLAM_E_INVALID;
    }
    VS_CUDA(cudaFree(nullptr));                       // make the primary context current: cuvid runs in it
    {
        int cr = 0;
        if (!nvdec_engine_ok(cu, &cr)) {
            vs_set_error("vslam_nvdec_create: NVDEC does not take codec %d on this device (cuvidGetDecoderCaps: CUresult %d%s); "
                         "there is no CPU decode fallback in this library",
                         codec, cr, cr == 100 ? " = no device: the container lacks the `video` driver capability" : "");
            return VSLAM_E_CUDA;
        }
    }
    vslam_nvdec* d = new (std::nothrow) vslam_nvdec();
    if (!d) return VSLAM_E_NOMEM;
    d->codec = cu;
    CUVIDPARSERPARAMS pp;
    memset(&pp, 0, sizeof(pp));
    pp.CodecType = cu;
    pp.ulMaxNumDecodeSurfaces = 1;                    // the sequence callback returns the real count
    pp.ulMaxDisplayDelay = 0;                         // frames come out as soon as they can be displayed
    pp.pUserData = d;
    pp.pfnSequenceCallback = on_sequence;
    pp.pfnDecodePicture = on_decode;
    pp.pfnDisplayPicture = on_display;
    const CUresult_t rc = api.CreateVideoParser(&d->parser, &pp);
    if (rc != 0) {
        vs_set_error("vslam_nvdec_create: cuvidCreateVideoParser failed with CUresult %d", rc);
        delete d;
        return VSLAM_E_CUDA;
    }
    *out = d;
    return VSLAM_OK;
}

extern "C" int vslam_nvdec_destroy(vslam_nvdec* d)
{
    if (!d) return VSLAM_OK;
    CuvidApi& api = cuvid();
    if (d->parser) api.DestroyVideoParser(d->parser);
    if (d->decoder) api.DestroyDecoder(d->decoder);
    delete d;
    return VSLAM_OK;
}

extern "C" int vslam_nvdec_set_colour(vslam_nvdec* d, int cy, int cbu, int cgu, int cgv, int crv, int y_offset)
{
    VS_REQUIRE(d, "null argument");
    d->coef = {cy, cbu, cgu, cgv, crv, y_offset};
    return VSLAM_OK;
}

extern "C" int vslam_nvdec_frame_size(const vslam_nvdec* d, int* width, int* height)
{
    VS_REQUIRE(d && width && height, "null argument");
    *width = d->width; *height = d->height;
    return VSLAM_OK;
}

extern "C" int vslam_nvdec_decode(vslam_nvdec* d, const uint8_t* h_packet, size_t bytes, int end_of_stream, int mode,
                                  uint8_t* d_out, int out_pitch, size_t frame_stride, int max_frames, int* n_frames,
                                  void* stream)
{
    VS_REQUIRE(d && n_frames, "null argument");
    VS_REQUIRE(mode >= VSLAM_NVDEC_LUMA && mode <= VSLAM_NVDEC_NV12, "unknown output mode");
    VS_REQUIRE(h_packet || bytes == 0, "null packet");
    VS_REQUIRE(max_frames >= 0 && (max_frames == 0 || d_out), "null output");
    *n_frames = 0;
    d->d_out = d_out; d->out_pitch = out_pitch; d->frame_stride = frame_stride; d->max_frames = max_frames;
    d->mode = mode; d->n_out = 0; d->dropped = 0; d->stream = (cudaStream_t)stream; d->error = 0;
    CUVIDSOURCEDATAPACKET pkt;
    memset(&pkt, 0, sizeof(pkt));
    pkt.payload = h_packet;
    pkt.payload_size = bytes;
    pkt.flags = CUVID_PKT_ENDOFPICTURE | (end_of_stream ? CUVID_PKT_ENDOFSTREAM : 0);
    const CUresult_t rc = cuvid().ParseVideoData(d->parser, &pkt);
    d->d_out = nullptr;
    *n_frames = d->n_out;
    if (d->error) return d->error;
    if (rc != 0) {
        vs_set_error("vslam_nvdec_decode: cuvidParseVideoData failed with CUresult %d", rc);
        return VSLAM_E_CUDA;
    }
    if (d->dropped) {
        vs_set_error("vslam_nvdec_decode: %d decoded frame(s) did not fit into max_frames = %d", d->dropped, max_frames);
        return VSLAM_E_CAPACITY;
    }
    return VSLAM_OK;
}
