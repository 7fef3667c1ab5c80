// Frame pre-processing on the device: u8 BGR HWC frame -> (x/256 - 0.5) -> bicubic resize by `scale`, zero padding
// -> fp32 CHW net input, one launch, one read of the frame, one write of the net input.
//
// Replaces the GPU path of op::CvMatToOpInput::createArray (reference: src/openpose/core/cvMatToOpInput.cpp:96-123):
// reorderAndNormalize (src/openpose/gpu/cuda.cu:11-25: HWC u8 -> CHW float, v * (1/256) - 0.5) followed by
// resizeAndPadRbgGpu (src/openpose/net/resizeAndMergeBase.cu:56-78,531-551: xs = (x + 0.5) / scale - 0.5, clamp-to-
// edge Catmull-Rom of cuda.hu:91-144 inside [0, ws*scale) x [0, hs*scale), zero outside). The reference keeps the
// normalised full-size frame in a float scratch (12 bytes per source pixel written and read back); here the
// normalisation is applied to the 16 taps as they are loaded (exact: v/256 - 0.5 has no rounding for 8-bit v).
// Algorithmic bytes per frame: 3*hs*ws (u8 read) + 12*ht*wt (fp32 write).
//
// cropWarpKernel: the hand / face crop of HandExtractorCaffe::cropFrame (src/openpose/hand/handExtractorCaffe.cpp:44-76)
// and FaceExtractorCaffe::forwardPass (src/openpose/face/faceExtractorCaffe.cpp:133-149): cv::warpAffine(INTER_LINEAR |
// WARP_INVERSE_MAP, BORDER_CONSTANT 0) of the u8 frame followed by uCharCvMatToFloatPtr (utilities/openCv.cpp:57-130),
// every crop of a frame in one launch. OpenCV is a dependency of the reference, not part of it; the kernel restates
// its published fixed-point algorithm (imgwarp.cpp warpAffine + remapBilinear: coordinates in 1/1024 px from
// cvRound of double products, rounded to 1/32 px, integer bilinear weights, +half >> shift) and is pinned against
// cv2 4.13 through the oracle and tests/golden.
#include "../common.cuh"
#include "bicubic.cuh"

namespace opb
{
    constexpr int kPreThreads = 256;

    // u8 -> v * (1/256) - 0.5 (cuda.cu:23) without the quarter-rate I2F: 0x4B000000 | v is the float 2^23 + v, and
    // (2^23 + v) / 256 - 32768.5 is exact in one FMA (32768 + v/256 needs 23 bits), so the result is the same float.
    __device__ __forceinline__ float normU8(unsigned v)
    {
        return __fmaf_rn(__uint_as_float(0x4B000000u | v), 1.f / 256.f, -32768.5f);
    }

    __global__ void __launch_bounds__(kPreThreads) preprocessKernel(
        float* __restrict__ dst, const unsigned char* __restrict__ src, int n, int hs, int ws, int ht, int wt,
        float scale)
    {
        // flat pixel index: no idle lanes when the net width is not a multiple of the block (656 = 2.56 * 256)
        const unsigned pix = blockIdx.x * kPreThreads + threadIdx.x;
        const size_t plane = (size_t)ht * wt;
        if (pix >= plane)
            return;
        const int y = pix / (unsigned)wt, x = pix - y * wt, f = blockIdx.y;
        float* o = dst + (size_t)f * 3 * plane + (size_t)y * wt + x;
        // reference :66: x < widthSource * rescaleFactor (float compare)
        if (!((float)x < mul((float)ws, scale) && (float)y < mul((float)hs, scale)))
        {
            o[0] = 0.f;
            o[plane] = 0.f;
            o[2 * plane] = 0.f;
            return;
        }
        int xi[4], yi[4];
        float dx, dy;
        taps(xi, dx, srcCoord(x, scale), ws);
        taps(yi, dy, srcCoord(y, scale), hs);
        const unsigned char* frame = src + (size_t)f * hs * ws * 3;
        const uintptr_t lo = (reinterpret_cast<uintptr_t>(src) + 3) & ~uintptr_t(3);
        const uintptr_t hi = reinterpret_cast<uintptr_t>(src) + (size_t)n * hs * ws * 3;
        const bool adjacent = xi[3] == xi[0] + 3; // away from the clamped borders the 4 taps are 12 adjacent bytes
        float t[3][4];
#pragma unroll
        for (int j = 0; j < 4; j++)
        {
            const unsigned char* row = frame + (size_t)yi[j] * ws * 3;
            float v[3][4];
            const uintptr_t p = reinterpret_cast<uintptr_t>(row + xi[0] * 3);
            const uintptr_t a = p & ~uintptr_t(3);
            if (adjacent && a >= lo && a + 16 <= hi)
            {
                // four aligned words cover the 12 bytes; PRMT drops each byte straight into the 2^23 + v float
                const unsigned* wp = reinterpret_cast<const unsigned*>(a);
                const unsigned sh = (unsigned)(p & 3) * 8;
                const unsigned w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2), w3 = __ldg(wp + 3);
                const unsigned u[3] = {__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh),
                                       __funnelshift_r(w2, w3, sh)};
#pragma unroll
                for (int k = 0; k < 12; k++)
                    v[k % 3][k / 3] = __fmaf_rn(__uint_as_float(__byte_perm(u[k / 4], 0x4B000000u, 0x7440 | (k & 3))),
                                                1.f / 256.f, -32768.5f);
            }
            else
            {
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int c = 0; c < 3; c++)
                        v[c][i] = normU8(row[xi[i] * 3 + c]);
            }
#pragma unroll
            for (int c = 0; c < 3; c++)
                t[c][j] = cubic1d(v[c][0], v[c][1], v[c][2], v[c][3], dx);
        }
#pragma unroll
        for (int c = 0; c < 3; c++)
            o[c * plane] = cubic1d(t[c][0], t[c][1], t[c][2], t[c][3], dy);
    }

    // One thread per crop pixel. m = the 2x3 inverse map (crop pixel -> frame position), doubles as cv::Mat CV_64F.
    __global__ void __launch_bounds__(kPreThreads) cropWarpKernel(
        float* __restrict__ dst, const unsigned char* __restrict__ frame, const double* __restrict__ mats, int hs,
        int ws, int ht, int wt)
    {
        const unsigned pix = blockIdx.x * kPreThreads + threadIdx.x;
        const size_t plane = (size_t)ht * wt;
        if (pix >= plane)
            return;
        const int y = pix / (unsigned)wt, x = pix - y * wt, f = blockIdx.y;
        const double* m = mats + 6 * f;
        // imgwarp.cpp: AB_BITS = 10, INTER_BITS = 5; adelta/bdelta per column, X0/Y0 per row, all cvRound (to even)
        const int adelta = __double2int_rn(__dmul_rn(__dmul_rn(m[0], (double)x), 1024.0));
        const int bdelta = __double2int_rn(__dmul_rn(__dmul_rn(m[3], (double)x), 1024.0));
        const int X0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(m[1], (double)y), m[2]), 1024.0)) + 16;
        const int Y0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(m[4], (double)y), m[5]), 1024.0)) + 16;
        const int X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5;
        // remap stores the integer part as short (saturate_cast<short>)
        const int sx = max(-32768, min(32767, X >> 5)), sy = max(-32768, min(32767, Y >> 5));
        const int fx = X & 31, fy = Y & 31;
        const bool x0in = (unsigned)sx < (unsigned)ws, x1in = (unsigned)(sx + 1) < (unsigned)ws;
        const bool y0in = (unsigned)sy < (unsigned)hs, y1in = (unsigned)(sy + 1) < (unsigned)hs;
        unsigned p[4][3];
#pragma unroll
        for (int k = 0; k < 4; k++)
        {
            const bool in = ((k & 1) ? x1in : x0in) && ((k & 2) ? y1in : y0in);
            const unsigned char* q = frame + ((size_t)(sy + (k >> 1)) * ws + (sx + (k & 1))) * 3;
#pragma unroll
            for (int c = 0; c < 3; c++)
                p[k][c] = in ? q[c] : 0u; // BORDER_CONSTANT, borderValue 0
        }
        // weights (32-fy)(32-fx), ... times 32 = the int16 table entries (exact, sum 32768); (sum + 2^14) >> 15
        const unsigned w00 = (32 - fy) * (32 - fx), w01 = (32 - fy) * fx, w10 = fy * (32 - fx), w11 = fy * fx;
        float* o = dst + (size_t)f * 3 * plane + pix;
#pragma unroll
        for (int c = 0; c < 3; c++)
        {
            const unsigned v = (w00 * p[0][c] + w01 * p[1][c] + w10 * p[2][c] + w11 * p[3][c] + 512u) >> 10;
            o[c * plane] = normU8(v);
        }
    }
}

extern "C" opb200_status opb200_preprocess(float* dst, const unsigned char* src, int n_frames, int src_h, int src_w,
                                           int dst_h, int dst_w, float scale, void* stream)
{
    using namespace opb;
    OPB_REQUIRE(dst && src, "preprocess: null argument");
    // reference: cvMatToOpInput.cpp:69-70
    OPB_REQUIRE(n_frames > 0 && src_h > 0 && src_w > 0, "Wrong input element (empty inputData).");
    OPB_REQUIRE(dst_h > 0 && dst_w > 0 && scale > 0.f, "preprocess: empty net input size or non-positive scale");
    OPB_REQUIRE(n_frames <= 65535 && (size_t)dst_h * dst_w < (size_t)1 << 31,
                "preprocess: too many frames / pixels for one launch");
    const dim3 grid((unsigned)(((size_t)dst_h * dst_w + kPreThreads - 1) / kPreThreads), n_frames);
    preprocessKernel<<<grid, kPreThreads, 0, static_cast<cudaStream_t>(stream)>>>(dst, src, n_frames, src_h, src_w,
                                                                                 dst_h, dst_w, scale);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}

extern "C" opb200_status opb200_crop_warp(float* dst, const unsigned char* frame, const double* matrices, int n_crops,
                                          int src_h, int src_w, int dst_h, int dst_w, void* stream)
{
    using namespace opb;
    OPB_REQUIRE(dst && frame && matrices, "crop_warp: null argument");
    OPB_REQUIRE(src_h > 0 && src_w > 0, "Empty cvInputData."); // handExtractorCaffe.cpp / faceExtractorCaffe.cpp checks
    OPB_REQUIRE(src_h < 32768 && src_w < 32768, "crop_warp: frame too large (remap coordinates are 16-bit)");
    OPB_REQUIRE(n_crops > 0 && n_crops <= 65535 && dst_h > 0 && dst_w > 0 &&
                    (size_t)dst_h * dst_w < (size_t)1 << 31,
                "crop_warp: wrong crop count or size");
    const dim3 grid((unsigned)(((size_t)dst_h * dst_w + kPreThreads - 1) / kPreThreads), n_crops);
    cropWarpKernel<<<grid, kPreThreads, 0, static_cast<cudaStream_t>(stream)>>>(dst, frame, matrices, src_h, src_w,
                                                                               dst_h, dst_w);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}
