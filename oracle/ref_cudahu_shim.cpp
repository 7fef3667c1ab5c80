// TEST INFRASTRUCTURE (oracle) -- never linked into the product.
//
// CPU-side pin of the bicubic arithmetic: the reference's own device header
//   /root/reference/include/openpose_private/gpu/cuda.hu:91-186
// compiled UNMODIFIED as plain host C++ (outside nvcc `__device__` expands to nothing) with -ffp-contract=off, i.e.
// the operation order as written, one rounding per operation. The loops below only walk the pixels the way the
// reference's kernels index them (resizeAndMergeBase.cu:106-196: resize8TimesKernel incl. its 5x5 shared tile and
// first-block special case, resizeAndAddAndAverageKernel; :56-78 resizeAndPadKernel); every interpolated value comes
// from the header's bicubicInterpolate / bicubicInterpolate8Times. Linked into oracle/_ref/libopref_post.so.
// The reference's real CUDA build contracts these expressions into FMAs (nvcc default), so the GPU tests compare with
// oracle/_ref/libopref_cuda.so as well; this file is what the C restatement (post_oracle.c) is held bit-equal to.
#include <cmath>
using std::floor;
#include <openpose_private/gpu/cuda.hu>

extern "C"
{
    // resize8TimesKernel (resizeAndMergeBase.cu:106-162): block = (factor, factor) threads, one source pixel per block.
    void refhu_resize8(float* target, const float* source, int channels, int heightSource, int widthSource,
                       int heightTarget, int widthTarget, int rescaleFactor)
    {
        const int f = rescaleFactor;
        for (int c = 0; c < channels; c++)
            for (int y = 0; y < heightTarget; y++)
                for (int x = 0; x < widthTarget; x++)
                {
                    const int blockX = x / f, blockY = y / f, threadX = x % f, threadY = y % f;
                    const float xSource = (x + 0.5f) / float(f) - 0.5f;
                    const float ySource = (y + 0.5f) / float(f) - 0.5f;
                    float* out = target + ((size_t)c * heightTarget + y) * widthTarget + x;
                    if (blockX < 1 || blockY < 1)
                    {
                        *out = op::bicubicInterpolate(source + (size_t)c * heightSource * widthSource, xSource, ySource,
                                                      widthSource, heightSource, widthSource);
                        continue;
                    }
                    float tile[25];
                    const float minSourceXFloat = (blockX * f + 0.5f) / float(f) - 0.5f;
                    const int minSourceXInt = int(floor(minSourceXFloat)) - 1;
                    const float minSourceYFloat = (blockY * f + 0.5f) / float(f) - 0.5f;
                    const int minSourceYInt = int(floor(minSourceYFloat)) - 1;
                    for (int id = 0; id < 25; id++)
                    {
                        const int xClean = op::fastTruncateCuda(minSourceXInt + id % 5, 0, widthSource - 1);
                        const int yClean = op::fastTruncateCuda(minSourceYInt + id / 5, 0, heightSource - 1);
                        tile[id] = source[((size_t)c * heightSource + yClean) * widthSource + xClean];
                    }
                    *out = op::bicubicInterpolate8Times(tile, xSource, ySource, widthSource, heightSource, threadX,
                                                        threadY);
                }
    }

    // resizeAndAddAndAverageKernel (resizeAndMergeBase.cu:165-196).
    void refhu_resize_average(float* target, const float* const* sources, const int* heightSources,
                              const int* widthSources, const float* scaleHeights, const float* scaleWidths,
                              int counter, int channels, int heightTarget, int widthTarget)
    {
        for (int c = 0; c < channels; c++)
            for (int y = 0; y < heightTarget; y++)
                for (int x = 0; x < widthTarget; x++)
                {
                    float interpolated = 0.f;
                    for (int i = 0; i < counter; i++)
                    {
                        const float xSource = (x + 0.5f) / scaleWidths[i] - 0.5f;
                        const float ySource = (y + 0.5f) / scaleHeights[i] - 0.5f;
                        interpolated += op::bicubicInterpolate(
                            sources[i] + (size_t)c * heightSources[i] * widthSources[i], xSource, ySource,
                            widthSources[i], heightSources[i], widthSources[i]);
                    }
                    target[((size_t)c * heightTarget + y) * widthTarget + x] = interpolated / float(counter);
                }
    }

    // resizeAndPadKernel, float source (resizeAndMergeBase.cu:56-78): 3 CHW planes, zero outside the scaled image.
    void refhu_resize_and_pad(float* target, const float* source, int heightSource, int widthSource,
                              int heightTarget, int widthTarget, float rescaleFactor)
    {
        for (int c = 0; c < 3; c++)
            for (int y = 0; y < heightTarget; y++)
                for (int x = 0; x < widthTarget; x++)
                {
                    float* out = target + ((size_t)c * heightTarget + y) * widthTarget + x;
                    if (x < widthSource * rescaleFactor && y < heightSource * rescaleFactor)
                    {
                        const float xSource = (x + 0.5f) / rescaleFactor - 0.5f;
                        const float ySource = (y + 0.5f) / rescaleFactor - 0.5f;
                        *out = op::bicubicInterpolate(source + (size_t)c * heightSource * widthSource, xSource,
                                                      ySource, widthSource, heightSource, widthSource);
                    }
                    else
                        *out = 0;
                }
    }
}
