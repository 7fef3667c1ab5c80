/* TEST INFRASTRUCTURE (oracle) -- plain-C CPU restatement of the reference's post-processing.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library. It is the CHECKER for openpose_b200/csrc; it is never measured as the product and
 * never shipped. Built by oracle/Makefile with -O3 -ffp-contract=off (no FMA contraction, like the
 * reference's default x86-64 build, /root/reference/CMakeLists.txt:100-137).
 *
 * Pinning: the reference ships no golden vectors for this path. Each function
 * here is pinned instead against the reference's OWN object code (oracle/_ref/libopref_post.so,
 * built from /root/reference sources) by tests/test_oracle_vs_ref.py, and against the committed
 * fixtures under tests/golden/ that were generated from that object code.
 * The CUDA-build semantics (opo_nms / opo_pair_scores mode 1, opo_resize_merge / opo_preprocess with contract = 1)
 * restate what the reference's CUDA kernels compute INCLUDING the FMA contraction nvcc applies to them by default
 * (spelled with fmaf(), which is exactly fused whatever the host ISA); they are pinned on the GPU box against the
 * reference's own kernels (oracle/_ref/libopref_cuda.so, tests/test_refcuda_gpu.py). The as-written bicubic
 * (contract = 0) is pinned on the CPU against include/openpose_private/gpu/cuda.hu compiled as host C++
 * (oracle/ref_cudahu_shim.cpp, tests/test_oracle_vs_ref.py).
 *
 * File references below are relative to /root/reference.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define OPO_MODE_CPU 0  /* semantics of the reference's CPU_ONLY build (the judged mode) */
#define OPO_MODE_CUDA 1 /* semantics of the reference's CUDA kernels as nvcc builds them (-fmad=true) */

/* include/openpose/utilities/fastMath.hpp:29-32 (positiveIntRound): add 0.5 then truncate. */
static int pos_int_round(float a) { return (int)(a + 0.5f); }
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* ------------------------------------------------------------------------------------------------
 * NMS. CPU semantics: src/openpose/net/nmsBase.cpp:7-68 (peak predicate), :71-107 (7x7 centroid),
 * :110-170 (raster scan, first maxPeaks-1... peaks). CUDA semantics: src/openpose/net/nmsBase.cu:64-89.
 * peaks layout: (channels, maxPeaks+1, 3); row 0 col 0 = count. Rows beyond count are left untouched.
 * ------------------------------------------------------------------------------------------------ */
static int is_peak_cpu(const float* s, int w, int h, float thr, int x, int y)
{
    const float v = s[y * w + x];
    if (1 < x && x < w - 2 && 1 < y && y < h - 2)
    {
        if (!(v > thr))
            return 0;
        for (int dy = -1; dy <= 1; dy++)
            for (int dx = -1; dx <= 1; dx++)
                if ((dx || dy) && !(v > s[(y + dy) * w + x + dx]))
                    return 0;
        return 1;
    }
    if (x == 1 || x == w - 2 || y == 1 || y == h - 2)
    {
        if (!(v > thr))
            return 0;
        for (int dy = -1; dy <= 1; dy++)
            for (int dx = -1; dx <= 1; dx++)
            {
                if (!dx && !dy)
                    continue;
                const int xx = x + dx, yy = y + dy;
                const float nb = (xx >= 0 && xx < w && yy >= 0 && yy < h) ? s[yy * w + xx] : thr;
                if (!(v >= nb))
                    return 0;
            }
        return 1;
    }
    return 0;
}

static int is_peak_cuda(const float* s, int w, int h, float thr, int x, int y)
{
    if (!(0 < x && x < w - 1 && 0 < y && y < h - 1))
        return 0;
    const float v = s[y * w + x];
    if (!(v > thr))
        return 0;
    for (int dy = -1; dy <= 1; dy++)
        for (int dx = -1; dx <= 1; dx++)
            if ((dx || dy) && !(v > s[(y + dy) * w + x + dx]))
                return 0;
    return 1;
}

/* mode 1: nvcc fuses xAcc += x*score / yAcc += y*score of writeResultKernel (nmsBase.cu:211-213) into FFMAs
 * (SASS of writeResultKernel<float>: two FFMA + one FADD per tap). */
static void refine_peak(float* out, const float* s, int px, int py, int w, int h, float offx, float offy, int mode)
{
    float xAcc = 0.f, yAcc = 0.f, sAcc = 0.f;
    for (int dy = -3; dy <= 3; dy++)
    {
        const int y = py + dy;
        if (y < 0 || y >= h)
            continue;
        for (int dx = -3; dx <= 3; dx++)
        {
            const int x = px + dx;
            if (x < 0 || x >= w)
                continue;
            const float sc = s[y * w + x];
            if (sc > 0)
            {
                if (mode == OPO_MODE_CUDA)
                {
                    xAcc = fmaf((float)x, sc, xAcc);
                    yAcc = fmaf((float)y, sc, yAcc);
                }
                else
                {
                    xAcc += (float)x * sc;
                    yAcc += (float)y * sc;
                }
                sAcc += sc;
            }
        }
    }
    out[0] = xAcc / sAcc + offx;
    out[1] = yAcc / sAcc + offy;
    out[2] = s[py * w + px];
}

/* Returns 0, or 1 on an invalid threshold (nmsBase.cpp:121-122). */
int opo_nms(float* peaks, const float* heat, float thr, int channels, int h, int w, int maxPeaks,
            float offx, float offy, int mode)
{
    if (thr < 0 || thr > 1.0)
        return 1;
    for (int c = 0; c < channels; c++)
    {
        const float* s = heat + (size_t)c * h * w;
        float* t = peaks + (size_t)c * (maxPeaks + 1) * 3;
        int count = 0;
        for (int y = 0; y < h && count < maxPeaks; y++)
            for (int x = 0; x < w && count < maxPeaks; x++)
            {
                const int pk = mode == OPO_MODE_CPU ? is_peak_cpu(s, w, h, thr, x, y) : is_peak_cuda(s, w, h, thr, x, y);
                if (pk)
                {
                    count++;
                    refine_peak(t + 3 * count, s, x, y, w, h, offx, offy, mode);
                }
            }
        t[0] = (float)count;
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * resizeAndMerge, CUDA semantics.
 * include/openpose_private/gpu/cuda.hu:91-144 (index clamp, Catmull-Rom a=-0.5, evaluation order),
 * src/openpose/net/resizeAndMergeBase.cu:106-162 (x8: xs=(x+0.5)/8-0.5) and :165-196 (multi-scale:
 * xs=(x+0.5)/scaleW_i-0.5, sum in scale order, divide by count).
 * contract = 0: cubicInterpolate as written, one rounding per operation (= cuda.hu compiled as host C++).
 * contract = 1: the FMA chain nvcc makes of it in the reference's CUDA build (SASS of resize8TimesKernel<float>,
 *   resizeAndAddAndAverageKernel<float>, resizeAndPadKernel<float>: A = fma(v3,.5,fma(v2,-1.5,fma(v1,1.5,v0*-.5))),
 *   B = fma(v3,-.5,fma(v2,2,fma(v1,-2.5,v0))), t = ((A*d)*d)*d, t = fma(B*d,d,t), t = fma((v0-v2)*-.5,d,t), v1 + t).
 * ------------------------------------------------------------------------------------------------ */
static float cubic1d(float v0, float v1, float v2, float v3, float d, int contract)
{
    if (contract)
    {
        const float A = fmaf(v3, 0.5f, fmaf(v2, -1.5f, fmaf(v1, 1.5f, v0 * -0.5f)));
        const float B = fmaf(v3, -0.5f, fmaf(v2, 2.f, fmaf(v1, -2.5f, v0)));
        float t = ((A * d) * d) * d;
        t = fmaf(B * d, d, t);
        t = fmaf((v0 - v2) * -0.5f, d, t);
        return v1 + t;
    }
    return (-0.5f * v0 + 1.5f * v1 - 1.5f * v2 + 0.5f * v3) * d * d * d
           + (v0 - 2.5f * v1 + 2.f * v2 - 0.5f * v3) * d * d
           - 0.5f * (v0 - v2) * d
           + v1;
}

static float bicubic_at(const float* src, float xs, float ys, int w, int h, int contract)
{
    int xi[4], yi[4];
    xi[1] = imin(w - 1, imax(0, (int)floorf(xs)));
    xi[0] = imax(0, xi[1] - 1);
    xi[2] = imin(w - 1, xi[1] + 1);
    xi[3] = imin(w - 1, xi[2] + 1);
    const float dx = xs - (float)xi[1];
    yi[1] = imin(h - 1, imax(0, (int)floorf(ys)));
    yi[0] = imax(0, yi[1] - 1);
    yi[2] = imin(h - 1, yi[1] + 1);
    yi[3] = imin(h - 1, yi[2] + 1);
    const float dy = ys - (float)yi[1];
    float t[4];
    for (int i = 0; i < 4; i++)
    {
        const float* r = src + (size_t)yi[i] * w;
        t[i] = cubic1d(r[xi[0]], r[xi[1]], r[xi[2]], r[xi[3]], dx, contract);
    }
    return cubic1d(t[0], t[1], t[2], t[3], dy, contract);
}

/* srcs[i]: (channels, hs[i], ws[i]); scaleW/H[i] as computed by resizeAndMergeBase.cu:379-380,475-477
 * (for one scale both are the integer ratio, e.g. 8). dst: (channels, ht, wt). */
void opo_resize_merge(float* dst, const float* const* srcs, const int* hs, const int* ws,
                      const float* scaleW, const float* scaleH, int nscales, int channels, int ht, int wt, int contract)
{
    for (int c = 0; c < channels; c++)
        for (int y = 0; y < ht; y++)
            for (int x = 0; x < wt; x++)
            {
                float acc = 0.f;
                float v = 0.f;
                for (int i = 0; i < nscales; i++)
                {
                    const float xs = ((float)x + 0.5f) / scaleW[i] - 0.5f;
                    const float ys = ((float)y + 0.5f) / scaleH[i] - 0.5f;
                    v = bicubic_at(srcs[i] + (size_t)c * hs[i] * ws[i], xs, ys, ws[i], hs[i], contract);
                    acc += v;
                }
                dst[((size_t)c * ht + y) * wt + x] = nscales == 1 ? v : acc / (float)nscales;
            }
}

/* ------------------------------------------------------------------------------------------------
 * Frame pre-processing, the reference's device path of CvMatToOpInput::createArray
 * (src/openpose/core/cvMatToOpInput.cpp:108-148): reorderAndNormalize (src/openpose/gpu/cuda.cu:11-25)
 * into a planar float scratch, then resizeAndPadKernel (src/openpose/net/resizeAndMergeBase.cu:56-78).
 * The reference only has device code for this; contract = 1 evaluates the bicubic as the reference's compiled kernel
 * does (see cubic1d), contract = 0 as written. (v * (1/256) - 0.5 is exact for 8-bit v either way.)
 * src: (hs, ws, 3) u8; dst: (3, ht, wt). Returns 0, or 1 when the scratch cannot be allocated.
 * ------------------------------------------------------------------------------------------------ */
int opo_preprocess(float* dst, const unsigned char* src, int hs, int ws, int ht, int wt, float scale, int contract)
{
    const size_t area = (size_t)hs * ws;
    float* planar = (float*)malloc(3 * area * sizeof(float));
    if (!planar)
        return 1;
    for (int c = 0; c < 3; c++)
        for (size_t i = 0; i < area; i++)
            planar[c * area + i] = (float)src[i * 3 + c] * (1 / 256.f) - 0.5f;
    for (int c = 0; c < 3; c++)
        for (int y = 0; y < ht; y++)
            for (int x = 0; x < wt; x++)
            {
                float v = 0.f;
                if ((float)x < (float)ws * scale && (float)y < (float)hs * scale)
                {
                    const float xs = ((float)x + 0.5f) / scale - 0.5f;
                    const float ys = ((float)y + 0.5f) / scale - 0.5f;
                    v = bicubic_at(planar + c * area, xs, ys, ws, hs, contract);
                }
                dst[((size_t)c * ht + y) * wt + x] = v;
            }
    free(planar);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Hand / face crop: cv::warpAffine(frame, M, size, INTER_LINEAR | WARP_INVERSE_MAP, BORDER_CONSTANT, 0) +
 * uCharCvMatToFloatPtr(normalize = 1), as called by src/openpose/hand/handExtractorCaffe.cpp:44-76 and
 * src/openpose/face/faceExtractorCaffe.cpp:133-149 (src/openpose/utilities/openCv.cpp:57-130: v/256 - 0.5, CHW).
 * cv::warpAffine lives in OpenCV (a dependency, not vendored in the reference; any 2.4.x-4.x the user installed).
 * This restates its published 8-bit bilinear algorithm (modules/imgproc/src/imgwarp.cpp: AB_BITS 10, INTER_BITS 5,
 * INTER_REMAP_COEF_BITS 15, cvRound = round half to even) and is pinned against cv2 4.13.0 (tests/test_crop_cpu.py,
 * tests/golden/crop_warp.npz). m: the 2x3 inverse map as doubles. frame (hs, ws, 3) u8 -> dst (3, ht, wt).
 * ------------------------------------------------------------------------------------------------ */
static int sat_short(int v) { return v < -32768 ? -32768 : (v > 32767 ? 32767 : v); }

/* the warp alone: frame (hs, ws, 3) u8 -> dst (ht, wt, 3) u8, what cv::warpAffine returns */
void opo_warp_affine_u8(unsigned char* dst, const unsigned char* frame, const double* m, int hs, int ws, int ht, int wt)
{
    for (int y = 0; y < ht; y++)
    {
        const int X0 = (int)lrint((m[1] * y + m[2]) * 1024) + 16;
        const int Y0 = (int)lrint((m[4] * y + m[5]) * 1024) + 16;
        for (int x = 0; x < wt; x++)
        {
            const int adelta = (int)lrint(m[0] * x * 1024), bdelta = (int)lrint(m[3] * x * 1024);
            const int X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5;
            const int sx = sat_short(X >> 5), sy = sat_short(Y >> 5), fx = X & 31, fy = Y & 31;
            /* the int16 weight table: saturate_cast<short>(w * 32768) of the float products, which are exact */
            const int w[4] = {(32 - fy) * (32 - fx) * 32, (32 - fy) * fx * 32, fy * (32 - fx) * 32, fy * fx * 32};
            for (int c = 0; c < 3; c++)
            {
                int acc = 0;
                for (int k = 0; k < 4; k++)
                {
                    const int xx = sx + (k & 1), yy = sy + (k >> 1);
                    if (xx >= 0 && xx < ws && yy >= 0 && yy < hs)
                        acc += w[k] * frame[((size_t)yy * ws + xx) * 3 + c];
                }
                int v = (acc + (1 << 14)) >> 15;
                v = v < 0 ? 0 : (v > 255 ? 255 : v);
                dst[((size_t)y * wt + x) * 3 + c] = (unsigned char)v;
            }
        }
    }
}

/* warp + uCharCvMatToFloatPtr(normalize = 1): (3, ht, wt) fp32 = v / 256 - 0.5 */
void opo_crop_warp(float* dst, const unsigned char* frame, const double* m, int hs, int ws, int ht, int wt)
{
    unsigned char* img = (unsigned char*)malloc((size_t)ht * wt * 3);
    if (!img)
        return;
    opo_warp_affine_u8(img, frame, m, hs, ws, ht, wt);
    for (int c = 0; c < 3; c++)
        for (int y = 0; y < ht; y++)
            for (int x = 0; x < wt; x++)
                dst[((size_t)c * ht + y) * wt + x] = (float)img[((size_t)y * wt + x) * 3 + c] * (1 / 256.f) - 0.5f;
    free(img);
}

/* ------------------------------------------------------------------------------------------------
 * getHeatMapsCopy: src/openpose/pose/poseExtractorNet.cpp:106-243 -- group by group: copy, then the scalar remap
 * loop for the ScaleMode (fastTruncate = compare-and-select of include/openpose/utilities/fastMath.hpp:72-88,
 * positiveIntRound = (int)(a + 0.5f)). heat: (channels, hw) of one frame = parts | background (has_bkg) | pafs.
 * mask bits: 1 parts, 2 background, 4 PAFs; mode = op::ScaleMode value. Returns the number of channels written,
 * -1 when the background is requested for a model without one (:196-200).
 * ------------------------------------------------------------------------------------------------ */
static float fast_max(float a, float b) { return a > b ? a : b; }
static float fast_min(float a, float b) { return a < b ? a : b; }
static float fast_truncate(float v, float lo, float hi) { return fast_min(hi, fast_max(lo, v)); }

int opo_heatmaps_export(float* dst, const float* heat, int hw, int parts, int has_bkg, int pafs, int mask, int mode)
{
    enum { ZeroToOne = 3, ZeroToOneFixedAspect, PlusMinusOne, PlusMinusOneFixedAspect, UnsignedChar, NoScale };
    size_t off = 0;
    int written = 0;
    for (int group = 0; group < 2; group++) /* body parts, then background: same rules (:130-194) */
    {
        if (!(mask & (1 << group)))
            continue;
        if (group == 1 && !has_bkg)
            return -1;
        const size_t vol = (size_t)(group == 0 ? parts : 1) * hw;
        float* d = dst + off;
        memcpy(d, heat + (group == 0 ? 0 : (size_t)parts * hw), vol * sizeof(float));
        if (mode != NoScale)
        {
            if (mode == PlusMinusOne || mode == PlusMinusOneFixedAspect)
                for (size_t i = 0; i < vol; i++)
                    d[i] = fast_truncate(d[i], 0.f, 1.f) * 2.f - 1.f;
            else if (mode == UnsignedChar)
                for (size_t i = 0; i < vol; i++)
                    d[i] = (float)pos_int_round(fast_truncate(d[i], 0.f, 1.f) * 255.f);
            else
                for (size_t i = 0; i < vol; i++)
                    d[i] = fast_truncate(d[i], 0.f, 1.f);
        }
        off += vol;
        written += group == 0 ? parts : 1;
    }
    if (mask & 4) /* PAFs (:203-229) */
    {
        const size_t vol = (size_t)pafs * hw;
        float* d = dst + off;
        memcpy(d, heat + (size_t)(parts + (has_bkg ? 1 : 0)) * hw, vol * sizeof(float));
        if (mode != NoScale)
        {
            if (mode == ZeroToOne || mode == ZeroToOneFixedAspect)
                for (size_t i = 0; i < vol; i++)
                    d[i] = fast_truncate(d[i], -1.f, 1.f) * 0.5f + 0.5f;
            else if (mode == UnsignedChar)
                for (size_t i = 0; i < vol; i++)
                    d[i] = (float)pos_int_round(fast_truncate(d[i], -1.f, 1.f) * 128.5f + 128.5f);
            else
                for (size_t i = 0; i < vol; i++)
                    d[i] = fast_truncate(d[i], -1.f, 1.f);
        }
        written += pafs;
    }
    return written;
}

/* ------------------------------------------------------------------------------------------------
 * maximum: src/openpose/net/maximumBase.cpp:8-57 (cv::minMaxLoc = first maximum in raster order),
 * src/openpose/net/maximumBase.cu:85-113. target (parts, 3) = (x, y, value); background (last) skipped.
 * ------------------------------------------------------------------------------------------------ */
void opo_maximum(float* target, const float* source, int parts, int h, int w)
{
    for (int p = 0; p < parts; p++)
    {
        const float* s = source + (size_t)p * h * w;
        int best = 0;
        for (int i = 1; i < h * w; i++)
            if (s[i] > s[best])
                best = i;
        target[3 * p] = (float)(best % w);
        target[3 * p + 1] = (float)(best / w);
        target[3 * p + 2] = s[best];
    }
}

/* ------------------------------------------------------------------------------------------------
 * Model tables (data restated from src/openpose/pose/poseParameters.cpp:253-262 POSE_MAP_INDEX and
 * :416-424 POSE_BODY_PART_PAIRS; checked entry by entry against oracle/_ref in
 * tests/test_oracle_vs_ref.py). model ids follow op::PoseModel (include/openpose/pose/enumClasses.hpp:9-30).
 * ------------------------------------------------------------------------------------------------ */
#define OPO_BODY_25 0
#define OPO_COCO_18 1

static const unsigned B25_PAIRS[] = {1,8, 1,2, 1,5, 2,3, 3,4, 5,6, 6,7, 8,9, 9,10, 10,11, 8,12, 12,13, 13,14,
                                     1,0, 0,15, 15,17, 0,16, 16,18, 2,17, 5,18, 14,19, 19,20, 14,21, 11,22,
                                     22,23, 11,24};
static const unsigned B25_MAP[] = {0,1, 14,15, 22,23, 16,17, 18,19, 24,25, 26,27, 6,7, 2,3, 4,5, 8,9, 10,11,
                                   12,13, 30,31, 32,33, 36,37, 34,35, 38,39, 20,21, 28,29, 40,41, 42,43,
                                   44,45, 46,47, 48,49, 50,51};
static const unsigned COCO_PAIRS[] = {1,2, 1,5, 2,3, 3,4, 5,6, 6,7, 1,8, 8,9, 9,10, 1,11, 11,12, 12,13, 1,0,
                                      0,14, 14,16, 0,15, 15,17, 2,16, 5,17};
static const unsigned COCO_MAP[] = {12,13, 20,21, 14,15, 16,17, 22,23, 24,25, 0,1, 2,3, 4,5, 6,7, 8,9, 10,11,
                                    28,29, 30,31, 34,35, 32,33, 36,37, 18,19, 26,27};

int opo_model_tables(int model, int* parts, int* pairs, const unsigned** pairTable, const unsigned** mapTable)
{
    if (model == OPO_BODY_25)
    {
        *parts = 25; *pairs = 26; *pairTable = B25_PAIRS; *mapTable = B25_MAP;
        return 0;
    }
    if (model == OPO_COCO_18)
    {
        *parts = 18; *pairs = 19; *pairTable = COCO_PAIRS; *mapTable = COCO_MAP;
        return 0;
    }
    return 1;
}

/* ------------------------------------------------------------------------------------------------
 * PAF line-integral score of one candidate pair.
 * mode 0: getScoreAB, src/openpose/net/bodyPartConnectorBase.cpp:12-75 (double compares, lower clamp,
 *         fall-through 0). mode 1: process(), src/openpose/net/bodyPartConnectorBase.cu:15-66 (no lower
 *         clamp, float sqrtf threshold, fall-through -1) with the contraction of the reference's CUDA build
 *         (SASS of pafScoreKernel<float>: norm^2 = fma(vx,vx,vy*vy), sample = fma(lm,step,s),
 *         score = fma(mapX,ux,mapY*uy)).
 * a, b point at (x, y, score) triplets.
 * ------------------------------------------------------------------------------------------------ */
float opo_paf_score(const float* a, const float* b, const float* mapX, const float* mapY, int w, int h,
                    float interThreshold, float interMinAbove, float nmsThreshold, int mode)
{
    const float vx = b[0] - a[0];
    const float vy = b[1] - a[1];
    const float vmax = fmaxf(fabsf(vx), fabsf(vy));
    const int n = imax(5, imin(25, pos_int_round(sqrtf(5 * vmax))));
    const float norm = mode == OPO_MODE_CUDA ? sqrtf(fmaf(vx, vx, vy * vy)) : sqrtf(vx * vx + vy * vy);
    if ((double)norm > 1e-6)
    {
        const float sx = a[0], sy = a[1];
        const float ux = vx / norm, uy = vy / norm;
        float sum = 0.f;
        unsigned count = 0;
        const float stepx = vx / (float)n, stepy = vy / (float)n;
        for (int lm = 0; lm < n; lm++)
        {
            int mx, my;
            if (mode == OPO_MODE_CPU)
            {
                mx = imax(0, imin(w - 1, pos_int_round(sx + (float)lm * stepx)));
                my = imax(0, imin(h - 1, pos_int_round(sy + (float)lm * stepy)));
            }
            else
            {
                mx = imin(w - 1, pos_int_round(fmaf((float)lm, stepx, sx)));
                my = imin(h - 1, pos_int_round(fmaf((float)lm, stepy, sy)));
            }
            const int idx = my * w + mx;
            const float sc = mode == OPO_MODE_CPU ? ux * mapX[idx] + uy * mapY[idx]
                                                  : fmaf(mapX[idx], ux, mapY[idx] * uy);
            if (sc > interThreshold)
            {
                sum += sc;
                count++;
            }
        }
        if ((float)count / (float)n > interMinAbove)
            return sum / (float)count;
        const float l2 = norm;
        if (mode == OPO_MODE_CPU)
        {
            const double thr = sqrt((double)(w * h)) / 150;
            if ((double)l2 < thr)
                return (float)((double)nmsThreshold + 1e-6);
        }
        else
        {
            const float thr = sqrtf((float)(w * h)) / 150;
            if (l2 < thr)
                return (float)((double)nmsThreshold + 1e-6);
        }
    }
    return mode == OPO_MODE_CPU ? 0.f : -1.f;
}

/* Score cube (pairs, maxPeaks, maxPeaks) as pafScoreKernel writes it (bodyPartConnectorBase.cu:108-145):
 * -1 where peakA >= nA or peakB >= nB. heat = full heat-map blob (channels, h, w). */
void opo_pair_scores(float* cube, const float* heat, const float* peaks, int model, int w, int h,
                     int maxPeaks, float interThreshold, float interMinAbove, float nmsThreshold, int mode)
{
    int parts, pairs;
    const unsigned *pt, *mt;
    if (opo_model_tables(model, &parts, &pairs, &pt, &mt))
        return;
    const size_t area = (size_t)w * h;
    const int po = 3 * (maxPeaks + 1);
    for (int p = 0; p < pairs; p++)
    {
        const float* A = peaks + pt[2 * p] * po;
        const float* B = peaks + pt[2 * p + 1] * po;
        const float* mapX = heat + (parts + 1 + mt[2 * p]) * area;
        const float* mapY = heat + (parts + 1 + mt[2 * p + 1]) * area;
        for (int i = 0; i < maxPeaks; i++)
            for (int j = 0; j < maxPeaks; j++)
            {
                float v = -1.f;
                if ((float)i < A[0] && (float)j < B[0])
                    v = opo_paf_score(A + 3 * (i + 1), B + 3 * (j + 1), mapX, mapY, w, h, interThreshold,
                                      interMinAbove, nmsThreshold, mode);
                cube[((size_t)p * maxPeaks + i) * maxPeaks + j] = v;
            }
    }
}

/* ------------------------------------------------------------------------------------------------
 * Person assembly.
 * people: growable list of rows; row[p] = flat index of the SCORE element of the chosen peak of part p
 * inside the peaks blob (part*3*(maxPeaks+1) + k*3 + 2), 0 = none. The reference keeps the part count in
 * row.back() (rows of limb 0 are two ints longer, bodyPartConnectorBase.cpp:384, which only moves where
 * back() lives); we keep it in .count.
 * ------------------------------------------------------------------------------------------------ */
#define OPO_MAX_PARTS 25
typedef struct { int row[OPO_MAX_PARTS]; int count; float score; int dead; } opo_person;
typedef struct { opo_person* v; int n, cap; } opo_people;

static opo_person* people_push(opo_people* P)
{
    if (P->n == P->cap)
    {
        P->cap = P->cap ? 2 * P->cap : 256;
        P->v = (opo_person*)realloc(P->v, sizeof(opo_person) * (size_t)P->cap);
    }
    opo_person* p = &P->v[P->n++];
    memset(p, 0, sizeof(*p));
    return p;
}

typedef struct { double score; int i, j; } opo_cand;
/* std::greater<std::tuple<double,int,int>> (bodyPartConnectorBase.cpp:346-348): descending lexicographic */
static int cand_desc(const void* pa, const void* pb)
{
    const opo_cand* a = (const opo_cand*)pa;
    const opo_cand* b = (const opo_cand*)pb;
    if (a->score != b->score) return a->score > b->score ? -1 : 1;
    if (a->i != b->i) return a->i > b->i ? -1 : 1;
    if (a->j != b->j) return a->j > b->j ? -1 : 1;
    return 0;
}

/* createPeopleVector, src/openpose/net/bodyPartConnectorBase.cpp:157-472 (BODY_25 / COCO_18 branches). */
static void create_people_cpu(opo_people* P, const float* heat, const float* peaks, int model, int w, int h,
                              int maxPeaks, float interThreshold, float interMinAbove, float nmsThreshold)
{
    int parts, pairs;
    const unsigned *pt, *mt;
    opo_model_tables(model, &parts, &pairs, &pt, &mt);
    const int po = 3 * (maxPeaks + 1);
    const size_t area = (size_t)w * h;
    opo_cand* cands = (opo_cand*)malloc(sizeof(opo_cand) * (size_t)maxPeaks * maxPeaks);
    int* conA = (int*)malloc(sizeof(int) * (size_t)maxPeaks);
    int* conB = (int*)malloc(sizeof(int) * (size_t)maxPeaks);
    float* conS = (float*)malloc(sizeof(float) * (size_t)maxPeaks);
    char* usedA = (char*)malloc((size_t)maxPeaks + 1);
    char* usedB = (char*)malloc((size_t)maxPeaks + 1);
    const int earFirst = parts == 18 ? 17 : 18; /* :400-405 */
    for (int pair = 0; pair < pairs; pair++)
    {
        const int partA = (int)pt[2 * pair], partB = (int)pt[2 * pair + 1];
        const float* A = peaks + partA * po;
        const float* B = peaks + partB * po;
        const int nA = pos_int_round(A[0]), nB = pos_int_round(B[0]);
        if (nA == 0 || nB == 0)
        {
            /* :190-289 (non-MPI): singletons of the non-empty side unless some person already holds them */
            const int part = nA == 0 ? partB : partA;
            const float* C = nA == 0 ? B : A;
            const int n = nA == 0 ? nB : nA;
            for (int i = 1; i <= n; i++)
            {
                const int off = part * po + i * 3 + 2;
                int found = 0;
                for (int q = 0; q < P->n && !found; q++)
                    found = P->v[q].row[part] == off;
                if (!found)
                {
                    opo_person* np = people_push(P);
                    np->row[part] = off;
                    np->count = 1;
                    np->score = C[i * 3 + 2];
                }
            }
            continue;
        }
        const float* mapX = heat + (parts + 1 + mt[2 * pair]) * area;
        const float* mapY = heat + (parts + 1 + mt[2 * pair + 1]) * area;
        int nc = 0;
        for (int i = 1; i <= nA; i++)
            for (int j = 1; j <= nB; j++)
            {
                const float s = opo_paf_score(A + 3 * i, B + 3 * j, mapX, mapY, w, h, interThreshold,
                                              interMinAbove, nmsThreshold, OPO_MODE_CPU);
                if ((double)s > 1e-6)
                {
                    cands[nc].score = (double)s;
                    cands[nc].i = i;
                    cands[nc].j = j;
                    nc++;
                }
            }
        qsort(cands, (size_t)nc, sizeof(opo_cand), cand_desc);
        /* greedy one-to-one, :350-372. The occupancy flags are set AFTER the counter==minAB break. */
        const int minAB = imin(nA, nB);
        memset(usedA, 0, (size_t)maxPeaks + 1);
        memset(usedB, 0, (size_t)maxPeaks + 1);
        int ncon = 0;
        for (int k = 0; k < nc; k++)
        {
            const int i = cands[k].i, j = cands[k].j;
            if (!usedA[i] && !usedB[j])
            {
                conA[ncon] = partA * po + i * 3 + 2;
                conB[ncon] = partB * po + j * 3 + 2;
                conS[ncon] = (float)cands[k].score;
                ncon++;
                if (ncon == minAB)
                    break;
                usedA[i] = 1;
                usedB[j] = 1;
            }
        }
        if (pair == 0)
        {
            for (int k = 0; k < ncon; k++) /* :378-393 */
            {
                opo_person* np = people_push(P);
                np->row[partA] = conA[k];
                np->row[partB] = conB[k];
                np->count = 2;
                np->score = peaks[conA[k]] + peaks[conB[k]] + conS[k];
            }
        }
        else if (pair == earFirst || pair == earFirst + 1)
        {
            for (int k = 0; k < ncon; k++) /* :399-428: fills gaps in EVERY matching person, no score/count */
                for (int q = 0; q < P->n; q++)
                {
                    int* row = P->v[q].row;
                    if (row[partA] == conA[k] && row[partB] == 0)
                        row[partB] = conB[k];
                    else if (row[partB] == conB[k] && row[partA] == 0)
                        row[partA] = conA[k];
                }
        }
        else
        {
            for (int k = 0; k < ncon; k++) /* :430-461 */
            {
                int found = 0;
                for (int q = 0; q < P->n; q++)
                    if (P->v[q].row[partA] == conA[k])
                    {
                        P->v[q].row[partB] = conB[k];
                        P->v[q].count++;
                        P->v[q].score += peaks[conB[k]] + conS[k];
                        found = 1;
                        break;
                    }
                if (!found)
                {
                    opo_person* np = people_push(P);
                    np->row[partA] = conA[k];
                    np->row[partB] = conB[k];
                    np->count = 2;
                    np->score = peaks[conA[k]] + peaks[conB[k]] + conS[k];
                }
            }
        }
    }
    free(cands); free(conA); free(conB); free(conS); free(usedA); free(usedB);
}

/* removePeopleBelowThresholdsAndFillFaces, :721-884 (parts < 135 => no face logic). Returns #valid. */
static int filter_people(int* valid, const opo_people* P, int parts, int minSubsetCnt, float minSubsetScore,
                         int maximizePositives)
{
    int n = 0;
    for (int q = 0; q < P->n; q++)
    {
        if (P->v[q].dead)
            continue;
        int counter = P->v[q].count;
        if (!maximizePositives && parts == 25)
        {
            int feet = 0;
            for (int i = 19; i < 25; i++)
                feet += P->v[q].row[i] > 0;
            const int before = counter;
            if (feet > 0)
                counter += 0 - feet; /* getKeypointCounter(...,19,25,0), :78-96 */
            if (counter != before && counter <= 4)
                continue;
        }
        if (counter >= minSubsetCnt && (P->v[q].score / (float)counter) >= minSubsetScore)
            valid[n++] = q;
    }
    if (n == 0 && !maximizePositives)
        return filter_people(valid, P, parts, minSubsetCnt, minSubsetScore, 1);
    return n;
}

/* peopleVectorToPeopleArray, :887-934. outRows (optional): (capacity, parts+1) ints = row + count. */
static void people_to_arrays(float* outKp, float* outScores, int* outRows, int capacity, const opo_people* P,
                             const int* valid, int nvalid, const float* peaks, int parts, int pairs, float scale)
{
    const float inv = 1 / (float)(parts + pairs);
    for (int k = 0; k < nvalid && k < capacity; k++)
    {
        const opo_person* p = &P->v[valid[k]];
        for (int b = 0; b < parts; b++)
        {
            float* o = outKp + ((size_t)k * parts + b) * 3;
            const int idx = p->row[b];
            if (idx > 0)
            {
                o[0] = peaks[idx - 2] * scale;
                o[1] = peaks[idx - 1] * scale;
                o[2] = peaks[idx];
            }
            else
                o[0] = o[1] = o[2] = 0.f;
            if (outRows)
                outRows[(size_t)k * (parts + 1) + b] = idx;
        }
        if (outRows)
            outRows[(size_t)k * (parts + 1) + parts] = p->count;
        outScores[k] = p->score * inv;
    }
}

/* connectBodyPartsCpu, :1328-1377. Returns the number of people (may exceed capacity; only capacity written),
 * -1 for an unsupported model. */
int opo_connect_cpu(const float* heat, const float* peaks, int model, int w, int h, int maxPeaks,
                    float interMinAbove, float interThreshold, int minSubsetCnt, float minSubsetScore,
                    float nmsThreshold, float scale, int maximizePositives, float* outKp, float* outScores,
                    int* outRows, int capacity)
{
    int parts, pairs;
    const unsigned *pt, *mt;
    if (opo_model_tables(model, &parts, &pairs, &pt, &mt))
        return -1;
    opo_people P = {0, 0, 0};
    create_people_cpu(&P, heat, peaks, model, w, h, maxPeaks, interThreshold, interMinAbove, nmsThreshold);
    int* valid = (int*)malloc(sizeof(int) * (size_t)(P.n + 1));
    const int n = filter_people(valid, &P, parts, minSubsetCnt, minSubsetScore, maximizePositives);
    people_to_arrays(outKp, outScores, outRows, capacity, &P, valid, n, peaks, parts, pairs, scale);
    free(valid);
    free(P.v);
    return n;
}

/* ------------------------------------------------------------------------------------------------
 * CUDA-build host semantics (secondary mode): pafPtrIntoVector (:475-542) + pafVectorIntoPeopleVector
 * (:545-718), fed by the (pairs, maxPeaks, maxPeaks) score cube of opo_pair_scores.
 * ------------------------------------------------------------------------------------------------ */
typedef struct { float total, paf; int pair, ia, ib; } opo_conn;
/* std::greater<std::tuple<double,double,int,int,int>>: descending lexicographic */
static int conn_desc(const void* pa, const void* pb)
{
    const opo_conn* a = (const opo_conn*)pa;
    const opo_conn* b = (const opo_conn*)pb;
    if (a->total != b->total) return a->total > b->total ? -1 : 1;
    if (a->paf != b->paf) return a->paf > b->paf ? -1 : 1;
    if (a->pair != b->pair) return a->pair > b->pair ? -1 : 1;
    if (a->ia != b->ia) return a->ia > b->ia ? -1 : 1;
    if (a->ib != b->ib) return a->ib > b->ib ? -1 : 1;
    return 0;
}

int opo_connect_from_pair_scores(const float* cube, const float* peaks, int model, int maxPeaks,
                                 int minSubsetCnt, float minSubsetScore, float scale, int maximizePositives,
                                 float* outKp, float* outScores, int* outRows, int capacity)
{
    int parts, pairs;
    const unsigned *pt, *mt;
    if (opo_model_tables(model, &parts, &pairs, &pt, &mt))
        return -1;
    const int po = 3 * (maxPeaks + 1);
    opo_conn* conns = (opo_conn*)malloc(sizeof(opo_conn) * (size_t)pairs * maxPeaks * maxPeaks);
    int nc = 0;
    for (int pair = 0; pair < pairs; pair++)
    {
        const int partA = (int)pt[2 * pair], partB = (int)pt[2 * pair + 1];
        const int nA = pos_int_round(peaks[partA * po]), nB = pos_int_round(peaks[partB * po]);
        for (int a = 0; a < nA; a++)
            for (int b = 0; b < nB; b++)
            {
                const float s = cube[((size_t)pair * maxPeaks + a) * maxPeaks + b];
                if ((double)s > 1e-6)
                {
                    const float sA = peaks[partA * po + (a + 1) * 3 + 2];
                    const float sB = peaks[partB * po + (b + 1) * 3 + 2];
                    conns[nc].total = s + 0.1f * sA + 0.1f * sB;
                    conns[nc].paf = s;
                    conns[nc].pair = pair;
                    conns[nc].ia = a + 1;
                    conns[nc].ib = b + 1;
                    nc++;
                }
            }
    }
    qsort(conns, (size_t)nc, sizeof(opo_conn), conn_desc);

    opo_people P = {0, 0, 0};
    int* assigned = (int*)malloc(sizeof(int) * (size_t)parts * maxPeaks);
    for (int i = 0; i < parts * maxPeaks; i++)
        assigned[i] = -1;
    for (int k = 0; k < nc; k++)
    {
        const int partA = (int)pt[2 * conns[k].pair], partB = (int)pt[2 * conns[k].pair + 1];
        const int idxA = (partA * (maxPeaks + 1) + conns[k].ia) * 3 + 2;
        const int idxB = (partB * (maxPeaks + 1) + conns[k].ib) * 3 + 2;
        int* aAs = &assigned[partA * maxPeaks + conns[k].ia - 1];
        int* bAs = &assigned[partB * maxPeaks + conns[k].ib - 1];
        const float paf = conns[k].paf;
        if (*aAs < 0 && *bAs < 0)
        {
            const int id = P.n;
            opo_person* np = people_push(&P);
            np->row[partA] = idxA;
            np->row[partB] = idxB;
            np->count = 2;
            np->score = peaks[idxA] + peaks[idxB] + paf;
            *aAs = id;
            *bAs = id;
        }
        else if ((*aAs >= 0) != (*bAs >= 0))
        {
            const int have = *aAs >= 0 ? *aAs : *bAs;
            int* other = *aAs >= 0 ? bAs : aAs;
            const int part2 = *aAs >= 0 ? partB : partA;
            const int idx2 = *aAs >= 0 ? idxB : idxA;
            opo_person* p = &P.v[have];
            if (p->row[part2] == 0)
            {
                p->row[part2] = idx2;
                p->count++;
                p->score += peaks[idx2] + paf;
                *other = have;
            }
        }
        else if (*aAs == *bAs)
            P.v[*aAs].score += paf;
        else
        {
            const int a1 = *aAs < *bAs ? *aAs : *bAs;
            const int a2 = *aAs < *bAs ? *bAs : *aAs;
            opo_person* p1 = &P.v[a1];
            const opo_person* p2 = &P.v[a2];
            int complementary = 1;
            for (int part = 0; part < parts && complementary; part++)
                if (p1->row[part] > 0 && p2->row[part] > 0)
                    complementary = 0;
            if (complementary)
            {
                for (int part = 0; part < parts; part++)
                    if (p1->row[part] == 0)
                        p1->row[part] = p2->row[part];
                p1->count += p2->count;
                p1->score += p2->score + paf;
                P.v[a2].dead = 1; /* erased after the loop in the reference (:707-709) */
                for (int i = 0; i < parts * maxPeaks; i++)
                    if (assigned[i] == a2)
                        assigned[i] = a1;
            }
        }
    }
    int* valid = (int*)malloc(sizeof(int) * (size_t)(P.n + 1));
    const int n = filter_people(valid, &P, parts, minSubsetCnt, minSubsetScore, maximizePositives);
    people_to_arrays(outKp, outScores, outRows, capacity, &P, valid, n, peaks, parts, pairs, scale);
    free(valid); free(assigned); free(conns); free(P.v);
    return n;
}
