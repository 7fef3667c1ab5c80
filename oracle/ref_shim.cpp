// TEST INFRASTRUCTURE (oracle) -- never linked into the product.
//
// Link shim that lets the reference's own, UNMODIFIED post-processing sources
//   /root/reference/src/openpose/net/nmsBase.cpp
//   /root/reference/src/openpose/net/bodyPartConnectorBase.cpp
//   /root/reference/src/openpose/pose/poseParameters.cpp   (+ point, rectangle, errorAndLog, string)
// be built into oracle/_ref/libopref_post.so without OpenCV (see oracle/Makefile).
// It supplies
//   (1) the six op::Array<T> members those objects leave undefined (the reference defines them
//       in src/openpose/core/array.cpp:100-118,283-338,421-440,641-708, which needs OpenCV),
//   (2) op::Matrix's default constructor (src/openpose/core/matrix.cpp, needs OpenCV),
//   (3) op::getKeypointsRoi(Rectangle,Rectangle) (src/openpose/utilities/keypoint.cpp:587-640;
//       only reachable for >=135-part models, which are out of scope -> aborts),
//   (4) a plain C entry surface (ref_*) so Python (ctypes) can drive the reference functions,
//   (5) for the people post-steps (hand/handDetector.cpp, face/faceDetector.cpp, core/keepTopNPeople.cpp,
//       core/keypointScaler.cpp, utilities/keypoint.cpp -- all compiled unmodified; keypoint.cpp's renderer sees a
//       type-only opencv2/imgproc stand-in written by oracle/Makefile and is never called): the Array<T> copy / move /
//       slice constructors, clone, getVolume(a, b) of array.cpp:165-280,470-500 (storage plumbing only).
//   (6) for the output serialisation (filestream/fileStream.cpp, jsonOfstream.cpp, cocoJsonSaver.cpp, peopleJsonSaver.cpp,
//       fileSaver.cpp, utilities/fileSystem.cpp -- all compiled unmodified; fileStream.cpp's image / FileStorage functions
//       see a type-only opencv2/highgui stand-in and are never called): Matrix(const void*) / getConstCvMat (never run).
// Nothing here restates reference arithmetic: nms / PAF scoring / assembly all run in the
// reference's own object code.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <numeric>
#include <openpose/core/array.hpp>
#include <openpose/core/matrix.hpp>
#include <openpose/core/point.hpp>
#include <openpose/core/rectangle.hpp>
#include <openpose/core/keepTopNPeople.hpp>
#include <openpose/core/keypointScaler.hpp>
#include <openpose/face/faceDetector.hpp>
#include <openpose/filestream/cocoJsonSaver.hpp>
#include <openpose/filestream/peopleJsonSaver.hpp>
#include <openpose/hand/handDetector.hpp>
#include <openpose/net/bodyPartConnectorBase.hpp>
#include <openpose/net/nmsBase.hpp>
#include <openpose/pose/poseParameters.hpp>

namespace op
{
#ifndef OPB_REF_SHIM_NO_MATRIX   // (the hand / face harness links the reference's real core/matrix.cpp instead)
    // (2) Matrix(): an empty handle is all Array<T> needs.
    Matrix::Matrix() {}
    // (6) only referenced by fileStream.cpp's image / FileStorage helpers, which the oracle never calls
    Matrix::Matrix(const void*) { std::abort(); }
    const void* Matrix::getConstCvMat() const { std::abort(); }
#endif

    // (1) Array<T> members. Storage semantics: row-major, owned by a shared_ptr<T[]>.
    template<typename T>
    void Array<T>::reset(const std::vector<int>& sizes)
    {
        if (sizes.empty())
        {
            mSize.clear();
            mVolume = 0;
            spData.reset();
            pData = nullptr;
            return;
        }
        mSize = sizes;
        mVolume = 1;
        for (const auto s : sizes)
            mVolume *= (size_t)s;
        spData.reset(new T[mVolume], std::default_delete<T[]>());
        pData = spData.get();
    }

    template<typename T>
    void Array<T>::reset(const int size)
    {
        if (size > 0)
            reset(std::vector<int>{size});
        else
            reset(std::vector<int>{});
    }

    template<typename T>
    void Array<T>::reset(const std::vector<int>& sizes, const T value)
    {
        reset(sizes);
        for (size_t i = 0; i < mVolume; i++)
            pData[i] = value;
    }

    template<typename T>
    Array<T>::Array(const std::vector<int>& sizes)
    {
        reset(sizes);
    }

    template<typename T>
    Array<T>::Array(const std::vector<int>& sizes, const T value)   // array.cpp:126-136
    {
        reset(sizes, value);
    }

    template<typename T>
    int Array<T>::getSize(const int index) const
    {
        if ((unsigned)index < mSize.size())
            return mSize[index];
        return !mSize.empty();   // array.cpp:425-437 (Matlab style): 0 for an empty Array, 1 past the last dimension
    }

    template<typename T>
    T& Array<T>::commonAt(const int index) const
    {
        if (index < 0 || (size_t)index >= mVolume)
        {
            std::fprintf(stderr, "oracle shim: Array index %d out of bounds (volume %zu)\n", index, mVolume);
            std::abort();
        }
        return pData[index];
    }

    // (5) storage plumbing
    template<typename T>
    Array<T>::Array(const Array<T>& array) :
        mSize{array.mSize}, mVolume{array.mVolume}, spData{array.spData}, pData{array.pData}
    {
    }

    template<typename T>
    Array<T>::Array(Array<T>&& array) :
        mSize{array.mSize}, mVolume{array.mVolume}
    {
        std::swap(spData, array.spData);
        std::swap(pData, array.pData);
    }

    template<typename T>
    Array<T>::Array(const Array<T>& array, const int index, const bool noCopy)
    {
        if (array.getSize(0) <= index)
        {
            std::fprintf(stderr, "oracle shim: Array slice index out of range\n");
            std::abort();
        }
        auto sizes = array.getSize();
        sizes[0] = 1;
        reset(sizes);      // always the copying flavour: the alias flavour only saves the copy
        const auto area = mVolume;
        std::copy(array.pData + index * area, array.pData + (index + 1) * area, pData);
    }

    template<typename T>
    Array<T> Array<T>::clone() const
    {
        Array<T> array{mSize};
        std::copy(pData, pData + mVolume, array.pData);
        return array;
    }

    template<typename T>
    size_t Array<T>::getVolume(const int indexA, const int indexB) const
    {
        const auto indexBFinal = (indexB != -1 ? indexB : (int)mSize.size() - 1);
        if (indexA < 0 || indexA > indexBFinal || (size_t)indexBFinal >= mSize.size())
        {
            std::fprintf(stderr, "oracle shim: Array::getVolume indexes out of dimension\n");
            std::abort();
        }
        size_t v = 1;
        for (int i = indexA; i <= indexBFinal; i++)
            v *= (size_t)mSize[i];
        return v;
    }

    template<typename T>
    Matrix& Array<T>::getCvMat()
    {
        std::fprintf(stderr, "oracle shim: Array::getCvMat reached (rendering) - out of scope\n");
        std::abort();
    }

#ifndef OPB_REF_SHIM_NO_MATRIX
    void* Matrix::getCvMat()
    {
        std::fprintf(stderr, "oracle shim: Matrix::getCvMat reached (rendering) - out of scope\n");
        std::abort();
    }
#endif

#define OPREF_INSTANTIATE_ARRAY(T) \
    template Array<T>::Array(const Array<T>&, const int, const bool); \
    template Array<T> Array<T>::clone() const; \
    template size_t Array<T>::getVolume(const int, const int) const; \
    template Matrix& Array<T>::getCvMat(); \
    template Array<T>::Array(const std::vector<int>&); \
    template Array<T>::Array(const std::vector<int>&, const T); \
    template void Array<T>::reset(const int); \
    template void Array<T>::reset(const std::vector<int>&); \
    template void Array<T>::reset(const std::vector<int>&, const T); \
    template int Array<T>::getSize(const int) const; \
    template T& Array<T>::commonAt(const int) const
    OPREF_INSTANTIATE_ARRAY(float);
    OPREF_INSTANTIATE_ARRAY(double);

    // The copy / move constructors cannot be named in an explicit instantiation (the class also has a converting
    // constructor template); using them once emits them for the reference's objects to link against.
    template <typename T>
    size_t emitCopyAndMove(const Array<T>& a)
    {
        Array<T> copy{a};
        Array<T> moved{std::move(copy)};
        return moved.getVolume();
    }
    template size_t emitCopyAndMove(const Array<float>&);
    template size_t emitCopyAndMove(const Array<double>&);

    // (3)
    template <typename T>
    float getKeypointsRoi(const Rectangle<T>&, const Rectangle<T>&)
    {
        std::fprintf(stderr, "oracle shim: getKeypointsRoi reached (>=135-part model) - out of scope\n");
        std::abort();
    }
    template float getKeypointsRoi(const Rectangle<float>&, const Rectangle<float>&);
    template float getKeypointsRoi(const Rectangle<double>&, const Rectangle<double>&);
}

// (4) C entry surface -------------------------------------------------------------------------
static thread_local char gErr[1024];

extern "C"
{
    const char* ref_last_error() { return gErr; }

    // nmsCpu, src/openpose/net/nmsBase.cpp:110-170
    int ref_nms_cpu(float* target, int* kernel, const float* source, float threshold,
                    const int* targetSize, const int* sourceSize, float offX, float offY)
    {
        try
        {
            const std::array<int, 4> t{{targetSize[0], targetSize[1], targetSize[2], targetSize[3]}};
            const std::array<int, 4> s{{sourceSize[0], sourceSize[1], sourceSize[2], sourceSize[3]}};
            op::nmsCpu(target, kernel, source, threshold, t, s, op::Point<float>{offX, offY});
            return 0;
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return 1;
        }
    }

    static int copyOut(const op::Array<float>& kp, const op::Array<float>& sc, float* outKeypoints,
                       float* outScores, int capacityPeople, int parts)
    {
        const int people = kp.getSize(0);
        const int n = people < capacityPeople ? people : capacityPeople;
        if (n > 0)
        {
            std::memcpy(outKeypoints, kp.getConstPtr(), sizeof(float) * (size_t)n * parts * 3);
            std::memcpy(outScores, sc.getConstPtr(), sizeof(float) * (size_t)n);
        }
        return people;
    }

    // connectBodyPartsCpu, src/openpose/net/bodyPartConnectorBase.cpp:1328-1377. Returns #people (or -1).
    int ref_connect_cpu(const float* heatMap, const float* peaks, int poseModel, int width, int height,
                        int maxPeaks, float interMinAboveThreshold, float interThreshold, int minSubsetCnt,
                        float minSubsetScore, float defaultNmsThreshold, float scaleFactor,
                        int maximizePositives, float* outKeypoints, float* outScores, int capacityPeople)
    {
        try
        {
            op::Array<float> kp, sc;
            const auto model = (op::PoseModel)poseModel;
            op::connectBodyPartsCpu(kp, sc, heatMap, peaks, model, op::Point<int>{width, height}, maxPeaks,
                                    interMinAboveThreshold, interThreshold, minSubsetCnt, minSubsetScore,
                                    defaultNmsThreshold, scaleFactor, maximizePositives != 0);
            return copyOut(kp, sc, outKeypoints, outScores, capacityPeople, (int)op::getPoseNumberBodyParts(model));
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return -1;
        }
    }

    // Host half of connectBodyPartsGpu (bodyPartConnectorBase.cu:209-230): pafPtrIntoVector ->
    // pafVectorIntoPeopleVector -> removePeopleBelowThresholdsAndFillFaces -> peopleVectorToPeopleArray,
    // fed with a (pairs, maxPeaks, maxPeaks) score cube.
    int ref_connect_from_pair_scores(const float* pairScores, const float* peaks, int poseModel, int maxPeaks,
                                     int minSubsetCnt, float minSubsetScore, float scaleFactor,
                                     int maximizePositives, float* outKeypoints, float* outScores,
                                     int capacityPeople)
    {
        try
        {
            const auto model = (op::PoseModel)poseModel;
            const auto& pairs = op::getPosePartPairs(model);
            const auto parts = op::getPoseNumberBodyParts(model);
            const auto numberPairs = (unsigned int)(pairs.size() / 2);
            op::Array<float> cube({(int)numberPairs, maxPeaks, maxPeaks});
            std::memcpy(cube.getPtr(), pairScores, sizeof(float) * cube.getVolume());
            const auto connections = op::pafPtrIntoVector(cube, peaks, maxPeaks, pairs, numberPairs);
            auto people = op::pafVectorIntoPeopleVector(connections, peaks, maxPeaks, pairs, parts);
            int numberPeople = 0;
            std::vector<int> valid;
            op::removePeopleBelowThresholdsAndFillFaces(valid, numberPeople, people, parts, minSubsetCnt,
                                                        minSubsetScore, maximizePositives != 0, peaks);
            op::Array<float> kp, sc;
            op::peopleVectorToPeopleArray(kp, sc, scaleFactor, people, valid, peaks, numberPeople, parts,
                                          numberPairs);
            return copyOut(kp, sc, outKeypoints, outScores, capacityPeople, (int)parts);
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return -1;
        }
    }

    // Model tables, src/openpose/pose/poseParameters.cpp:253-367,416-537
    int ref_pose_number_body_parts(int poseModel) { return (int)op::getPoseNumberBodyParts((op::PoseModel)poseModel); }
    int ref_pose_add_bkg(int poseModel) { return op::addBkgChannel((op::PoseModel)poseModel) ? 1 : 0; }
    int ref_pose_max_peaks() { return (int)op::getPoseMaxPeaks(); }
    int ref_pose_part_pairs(int poseModel, unsigned int* out, int capacity)
    {
        const auto& v = op::getPosePartPairs((op::PoseModel)poseModel);
        for (size_t i = 0; i < v.size() && (int)i < capacity; i++)
            out[i] = v[i];
        return (int)v.size();
    }
    int ref_pose_map_index(int poseModel, unsigned int* out, int capacity)
    {
        const auto& v = op::getPoseMapIndex((op::PoseModel)poseModel);
        for (size_t i = 0; i < v.size() && (int)i < capacity; i++)
            out[i] = v[i];
        return (int)v.size();
    }
    // out = {nmsThreshold, interMinAbove, interThreshold, minSubsetCnt, minSubsetScore}
    void ref_pose_defaults(int poseModel, int maximizePositives, float* out)
    {
        const auto model = (op::PoseModel)poseModel;
        const bool mp = maximizePositives != 0;
        out[0] = op::getPoseDefaultNmsThreshold(model, mp);
        out[1] = op::getPoseDefaultConnectInterMinAboveThreshold(mp);
        out[2] = op::getPoseDefaultConnectInterThreshold(model, mp);
        out[3] = (float)op::getPoseDefaultMinSubsetCnt(mp);
        out[4] = op::getPoseDefaultConnectMinSubsetScore(mp);
    }

    // ---- people post-steps -------------------------------------------------------------------
    static op::Array<float> peopleArray(const float* kp, int people, int parts)
    {
        op::Array<float> a;
        if (people > 0)
        {
            a.reset({people, parts, 3});
            std::memcpy(a.getPtr(), kp, sizeof(float) * (size_t)people * parts * 3);
        }
        return a;
    }

    // HandDetector::detectHands, src/openpose/hand/handDetector.cpp:136-164. out: (people, 2, 4) = x, y, w, h.
    int ref_detect_hands(const float* kp, int people, int parts, int poseModel, float* out)
    {
        try
        {
            const op::HandDetector detector{(op::PoseModel)poseModel};
            const auto rects = detector.detectHands(peopleArray(kp, people, parts));
            for (size_t p = 0; p < rects.size(); p++)
                for (int h = 0; h < 2; h++)
                {
                    const auto& r = rects[p][h];
                    const float v[4] = {r.x, r.y, r.width, r.height};
                    std::memcpy(out + (p * 2 + h) * 4, v, sizeof(v));
                }
            return (int)rects.size();
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return -1;
        }
    }

    // FaceDetector::detectFaces, src/openpose/face/faceDetector.cpp:120-141. out: (people, 4).
    int ref_detect_faces(const float* kp, int people, int parts, int poseModel, float* out)
    {
        try
        {
            const op::FaceDetector detector{(op::PoseModel)poseModel};
            const auto rects = detector.detectFaces(peopleArray(kp, people, parts));
            for (size_t p = 0; p < rects.size(); p++)
            {
                const float v[4] = {rects[p].x, rects[p].y, rects[p].width, rects[p].height};
                std::memcpy(out + p * 4, v, sizeof(v));
            }
            return (int)rects.size();
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return -1;
        }
    }

    // KeepTopNPeople::keepTopPeople, src/openpose/core/keepTopNPeople.cpp:17-91. Returns the people kept.
    int ref_keep_top_people(const float* kp, const float* scores, int people, int parts, int numberPeopleMax,
                            float* out)
    {
        try
        {
            op::Array<float> sc;
            if (people > 0)
            {
                sc.reset(people);
                std::memcpy(sc.getPtr(), scores, sizeof(float) * people);
            }
            const op::KeepTopNPeople keeper{numberPeopleMax};
            const auto top = keeper.keepTopPeople(peopleArray(kp, people, parts), sc);
            if (!top.empty())
                std::memcpy(out, top.getConstPtr(), sizeof(float) * top.getVolume());
            return top.getSize(0);
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return -1;
        }
    }

    // KeypointScaler::scale, src/openpose/core/keypointScaler.cpp:77-104 (in place).
    int ref_scale_keypoints(float* kp, int people, int parts, int scaleMode, double scaleInputToOutput,
                            double scaleNetToOutput, int producerW, int producerH)
    {
        try
        {
            auto a = peopleArray(kp, people, parts);
            const op::KeypointScaler scaler{(op::ScaleMode)scaleMode};
            scaler.scale(a, scaleInputToOutput, scaleNetToOutput, op::Point<int>{producerW, producerH});
            if (!a.empty())
                std::memcpy(kp, a.getConstPtr(), sizeof(float) * a.getVolume());
            return 0;
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return -1;
        }
    }

    // PeopleJsonSaver::save, src/openpose/filestream/peopleJsonSaver.cpp:15-30 -> savePeopleJson, fileStream.cpp:306-344.
    // arrays: nArrays x (data, dims[3], nDims) with nDims in {0 (empty Array), 1, 3}; file = dir/fileName.json
    int ref_save_people_json(const char* dir, const char* fileName, int nArrays, const float* const* data,
                             const int* dims, const int* nDims, const char* const* names, const float* cand,
                             const int* candCounts, int nParts, int humanReadable)
    {
        try
        {
            std::vector<std::pair<op::Array<float>, std::string>> kv;
            for (int v = 0; v < nArrays; ++v)
            {
                op::Array<float> a;
                if (nDims[v] > 0)
                {
                    std::vector<int> sz(dims + 3 * v, dims + 3 * v + nDims[v]);
                    a.reset(sz);
                    if (a.getVolume() > 0)
                        std::memcpy(a.getPtr(), data[v], sizeof(float) * a.getVolume());
                }
                kv.emplace_back(a, names[v]);
            }
            std::vector<std::vector<std::array<float, 3>>> candidates((size_t)nParts);
            const float* p = cand;
            for (int part = 0; part < nParts; ++part)
                for (int c = 0; c < candCounts[part]; ++c, p += 3)
                    candidates[part].push_back(std::array<float, 3>{p[0], p[1], p[2]});
            const op::PeopleJsonSaver saver{dir};
            saver.save(kv, candidates, fileName, humanReadable != 0);
            return 0;
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return -1;
        }
    }

    // CocoJsonSaver (constructor, record per frame, destructor), src/openpose/filestream/cocoJsonSaver.cpp:28-277.
    // kp: frames concatenated, people[f] x parts x 3 each; scores likewise; names: one image name per frame.
    int ref_coco_json(const char* path, int poseModel, int humanReadable, int variants, int format, int variant,
                      int nFrames, const float* kp, const float* scores, const int* people, int parts,
                      const char* const* names)
    {
        try
        {
            op::CocoJsonSaver saver{path, (op::PoseModel)poseModel, humanReadable != 0, variants,
                                    (op::CocoJsonFormat)format, variant};
            for (int f = 0; f < nFrames; ++f)
            {
                op::Array<float> sc;
                if (people[f] > 0)
                {
                    sc.reset(people[f]);
                    std::memcpy(sc.getPtr(), scores, sizeof(float) * people[f]);
                }
                saver.record(peopleArray(kp, people[f], parts), sc, names[f], (unsigned long long)f);
                kp += (size_t)people[f] * parts * 3;
                scores += people[f];
            }
            return 0;
        }
        catch (const std::exception& e)
        {
            std::snprintf(gErr, sizeof(gErr), "%s", e.what());
            return -1;
        }
    }
}
