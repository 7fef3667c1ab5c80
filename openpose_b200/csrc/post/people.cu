// People post-steps on the extractor boundary, on the device, over the connector's output layout
// (keypoints (N,capacity,parts,3), scores (N,capacity), count (N)) so body -> rectangles -> crops -> hand / face
// needs no host round trip per frame:
//   keepTopKernel     op::KeepTopNPeople::keepTopPeople   src/openpose/core/keepTopNPeople.cpp:17-91
//   scaleKernel       op::KeypointScaler::scale           src/openpose/core/keypointScaler.cpp:7-104,
//                     op::scaleKeypoints2d                src/openpose/utilities/keypoint.cpp:104-168
//   rectanglesKernel  op::HandDetector::detectHands       src/openpose/hand/handDetector.cpp:9-46,136-164
//                     op::FaceDetector::detectFaces       src/openpose/face/faceDetector.cpp:22-141
// Helper arithmetic restated from utilities/keypoint.cpp: getDistance :11-25, getKeypointsRectangle :288-342,
// getKeypointsArea :377-388. Every float operation is a separate round-to-nearest op in the reference's order
// (the host build has no FMA contraction); sqrtf is IEEE like std::sqrt. The work is a few KB per frame: these
// kernels are latency-sized (one block per frame), they exist to keep the pipeline on the device.
#include "../common.cuh"

namespace opb
{
    constexpr int kPeopleThreads = 256;

    __device__ __forceinline__ float dist(const float* p, int a, int b)
    {
        const float dx = sub(p[a * 3], p[b * 3]), dy = sub(p[a * 3 + 1], p[b * 3 + 1]);
        return sqrtf(add(mul(dx, dx), mul(dy, dy)));
    }

    // getKeypointsArea(keypoints, person, threshold) = width * height of the box around the parts above threshold
    __device__ float keypointsArea(const float* p, int parts, float threshold)
    {
        float minX = 3.402823466e+38f, maxX = -3.402823466e+38f, minY = minX, maxY = maxX;
        for (int part = 0; part < parts; part++)
            if (p[3 * part + 2] > threshold)
            {
                const float x = p[3 * part], y = p[3 * part + 1];
                if (maxX < x) maxX = x;
                if (minX > x) minX = x;
                if (maxY < y) maxY = y;
                if (minY > y) minY = y;
            }
        if (maxX >= minX && maxY >= minY)
            return mul(sub(maxX, minX), sub(maxY, minY));
        return 0.f;
    }

    // One block per frame. final score = score * sqrt(area); threshold = the numberPeopleMax-th largest; people above
    // it are kept in order, people equal to it fill the remaining places in order (keepTopNPeople.cpp:33-75).
    __global__ void __launch_bounds__(kPeopleThreads) keepTopKernel(
        float* __restrict__ outKp, float* __restrict__ outScores, int* __restrict__ outCount,
        const float* __restrict__ kp, const float* __restrict__ scores, const int* __restrict__ count, int capacity,
        int parts, int nMax)
    {
        extern __shared__ float finalScore[]; // capacity floats
        __shared__ float sThreshold;
        __shared__ int sAbove;
        const int f = blockIdx.x;
        const int people = min(count[f], capacity);
        const float* fkp = kp + (size_t)f * capacity * parts * 3;
        float* okp = outKp + (size_t)f * capacity * parts * 3;
        const int area = parts * 3;
        if (!(people > nMax && nMax > 0))
        {
            // returned unchanged
            for (int i = threadIdx.x; i < people * area; i += blockDim.x)
                okp[i] = fkp[i];
            for (int i = threadIdx.x; i < people; i += blockDim.x)
                outScores[(size_t)f * capacity + i] = scores[(size_t)f * capacity + i];
            if (threadIdx.x == 0)
                outCount[f] = people;
            return;
        }
        for (int p = threadIdx.x; p < people; p += blockDim.x)
            finalScore[p] = mul(scores[(size_t)f * capacity + p], sqrtf(keypointsArea(fkp + (size_t)p * area, parts, 0.05f)));
        if (threadIdx.x == 0)
            sAbove = 0;
        __syncthreads();
        // threshold = sorted-descending[nMax - 1]: the value v with #(> v) < nMax <= #(>= v)
        for (int p = threadIdx.x; p < people; p += blockDim.x)
        {
            const float v = finalScore[p];
            int greater = 0, greaterEq = 0;
            for (int q = 0; q < people; q++)
            {
                greater += finalScore[q] > v;
                greaterEq += finalScore[q] >= v;
            }
            if (greater < nMax && nMax <= greaterEq)
            {
                sThreshold = v; // every thread that gets here holds the same value
                sAbove = greater;
            }
        }
        __syncthreads();
        const float threshold = sThreshold;
        const int onThresholdToAdd = nMax - sAbove;
        for (int p = threadIdx.x; p < people; p += blockDim.x)
        {
            const float v = finalScore[p];
            if (!(v >= threshold))
                continue;
            int aboveBefore = 0, equalBefore = 0;
            for (int q = 0; q < p; q++)
            {
                aboveBefore += finalScore[q] > threshold;
                equalBefore += finalScore[q] == threshold;
            }
            const bool keep = v > threshold || equalBefore + 1 <= onThresholdToAdd;
            if (keep)
            {
                const int slot = aboveBefore + min(equalBefore, onThresholdToAdd);
                for (int i = 0; i < area; i++)
                    okp[(size_t)slot * area + i] = fkp[(size_t)p * area + i];
                outScores[(size_t)f * capacity + slot] = scores[(size_t)f * capacity + p];
            }
        }
        if (threadIdx.x == 0)
            outCount[f] = nMax;
    }

    // scaleKeypoints2d, both overloads: x * scaleX [+ offsetX], y * scaleY [+ offsetY]; the score is untouched.
    __global__ void __launch_bounds__(kPeopleThreads) scaleKernel(
        float* __restrict__ kp, const int* __restrict__ count, int capacity, int parts, float sx, float sy, float ox,
        float oy, int withOffset)
    {
        const int f = blockIdx.y;
        const int n = min(count ? count[f] : capacity, capacity) * parts;
        const int i = blockIdx.x * kPeopleThreads + threadIdx.x;
        if (i >= n)
            return;
        float* p = kp + ((size_t)f * capacity * parts + i) * 3;
        if (withOffset)
        {
            p[0] = add(mul(p[0], sx), ox);
            p[1] = add(mul(p[1], sy), oy);
        }
        else
        {
            p[0] = mul(p[0], sx);
            p[1] = mul(p[1], sy);
        }
    }

    struct RectParts
    {
        int lWrist, lElbow, lShoulder, rWrist, rElbow, rShoulder; // HandDetector::mPoseIndexes
        int neck, nose, lEar, rEar, lEye, rEye;                   // FaceDetector members
    };

    __device__ void handRect(float* out, const float* p, int wrist, int elbow, int shoulder)
    {
        const float threshold = 0.03f, ratio = 0.33f; // handDetector.cpp:143,20
        float x = 0.f, y = 0.f, w = 0.f;
        if (p[wrist * 3 + 2] > threshold && p[elbow * 3 + 2] > threshold && p[shoulder * 3 + 2] > threshold)
        {
            x = add(p[wrist * 3], mul(ratio, sub(p[wrist * 3], p[elbow * 3])));
            y = add(p[wrist * 3 + 1], mul(ratio, sub(p[wrist * 3 + 1], p[elbow * 3 + 1])));
            const float dWE = dist(p, wrist, elbow), dES = dist(p, elbow, shoulder);
            const float scaled = mul(0.9f, dES);
            w = mul(1.5f, (dWE > scaled ? dWE : scaled));
        }
        out[0] = sub(x, divf(w, 2.f));
        out[1] = sub(y, divf(w, 2.f));
        out[2] = w;
        out[3] = w;
    }

    __device__ void faceRect(float* out, const float* p, const RectParts& i)
    {
        const float threshold = 0.25f; // faceDetector.cpp:127
        const bool neck = p[i.neck * 3 + 2] > threshold, nose = p[i.nose * 3 + 2] > threshold;
        const bool lEar = p[i.lEar * 3 + 2] > threshold, rEar = p[i.rEar * 3 + 2] > threshold;
        const bool lEye = p[i.lEye * 3 + 2] > threshold, rEye = p[i.rEye * 3 + 2] > threshold;
        float x = 0.f, y = 0.f, size = 0.f;
        int counter = 0;
        if (neck && nose)
        {
            if (lEye == lEar && rEye == rEar && lEye != rEye) // profile: one eye and its ear
            {
                const int eye = lEye ? i.lEye : i.rEye, ear = lEye ? i.lEar : i.rEar;
                x = add(x, divf(add(add(p[eye * 3], p[ear * 3]), p[i.nose * 3]), 3.f));
                y = add(y, divf(add(add(p[eye * 3 + 1], p[ear * 3 + 1]), p[i.nose * 3 + 1]), 3.f));
                size = add(size, mul(0.85f, add(add(dist(p, i.nose, eye), dist(p, i.nose, ear)), dist(p, i.neck, i.nose))));
            }
            else
            {
                x = add(x, divf(add(p[i.neck * 3], p[i.nose * 3]), 2.f));
                y = add(y, divf(add(p[i.neck * 3 + 1], p[i.nose * 3 + 1]), 2.f));
                size = add(size, mul(2.f, dist(p, i.neck, i.nose)));
            }
            counter++;
        }
        if (lEye && rEye)
        {
            x = add(x, divf(add(p[i.lEye * 3], p[i.rEye * 3]), 2.f));
            y = add(y, divf(add(p[i.lEye * 3 + 1], p[i.rEye * 3 + 1]), 2.f));
            size = add(size, mul(3.f, dist(p, i.lEye, i.rEye)));
            counter++;
        }
        if (lEar && rEar)
        {
            x = add(x, divf(add(p[i.lEar * 3], p[i.rEar * 3]), 2.f));
            y = add(y, divf(add(p[i.lEar * 3 + 1], p[i.rEar * 3 + 1]), 2.f));
            size = add(size, mul(2.f, dist(p, i.lEar, i.rEar)));
            counter++;
        }
        if (counter > 0)
        {
            x = divf(x, (float)counter);
            y = divf(y, (float)counter);
            size = divf(size, (float)counter);
        }
        out[0] = sub(x, divf(size, 2.f));
        out[1] = sub(y, divf(size, 2.f));
        out[2] = size;
        out[3] = size;
    }

    __global__ void __launch_bounds__(kPeopleThreads) rectanglesKernel(
        float* __restrict__ hands, float* __restrict__ faces, const float* __restrict__ kp,
        const int* __restrict__ count, int capacity, int parts, const RectParts idx)
    {
        const int f = blockIdx.y;
        const int person = blockIdx.x * kPeopleThreads + threadIdx.x;
        if (person >= min(count[f], capacity))
            return;
        const float* p = kp + ((size_t)f * capacity + person) * parts * 3;
        if (hands)
        {
            float* h = hands + ((size_t)f * capacity + person) * 8;
            handRect(h, p, idx.lWrist, idx.lElbow, idx.lShoulder);
            handRect(h + 4, p, idx.rWrist, idx.rElbow, idx.rShoulder);
        }
        if (faces)
            faceRect(faces + ((size_t)f * capacity + person) * 4, p, idx);
    }
}

extern "C" opb200_status opb200_keep_top_people(float* out_keypoints, float* out_scores, int* out_count,
                                                const float* keypoints, const float* scores, const int* count,
                                                int n_frames, int capacity, int parts, int number_people_max,
                                                void* stream)
{
    using namespace opb;
    OPB_REQUIRE(out_keypoints && out_scores && out_count && keypoints && scores && count, "keep_top_people: null argument");
    OPB_REQUIRE(n_frames > 0 && capacity > 0 && parts > 0, "keep_top_people: empty people arrays");
    OPB_REQUIRE((size_t)capacity * sizeof(float) <= 48 * 1024, "keep_top_people: capacity too large");
    keepTopKernel<<<n_frames, kPeopleThreads, capacity * sizeof(float), static_cast<cudaStream_t>(stream)>>>(
        out_keypoints, out_scores, out_count, keypoints, scores, count, capacity, parts, number_people_max);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}

extern "C" opb200_status opb200_keypoint_scale_and_offset(float out[4], int scale_mode, double scale_input_to_output,
                                                          double scale_net_to_output, int producer_w, int producer_h)
{
    // getScaleAndOffset, keypointScaler.cpp:7-42; out = offset x, offset y, scale x, scale y (the Rectangle fields)
    OPB_REQUIRE(out, "keypoint_scale_and_offset: null argument");
    const float w = (float)producer_w, h = (float)producer_h;
    const float big = (float)(producer_w > producer_h ? producer_w : producer_h);
    out[0] = out[1] = 0.f;
    switch (scale_mode)
    {
    case 2: out[2] = out[3] = float(scale_input_to_output); break;          // OutputResolution
    case 1: out[2] = out[3] = float(1. / scale_net_to_output); break;       // NetOutputResolution
    case 3: out[2] = 1.f / (w - 1.f); out[3] = 1.f / (h - 1.f); break;      // ZeroToOne
    case 4: out[2] = out[3] = 1.f / (big - 1.f); break;                     // ZeroToOneFixedAspect
    case 5: out[0] = out[1] = -1.f; out[2] = 2.f / (w - 1.f); out[3] = 2.f / (h - 1.f); break; // PlusMinusOne
    case 6: out[0] = out[1] = -1.f; out[2] = out[3] = 2.f / (big - 1.f); break;                // ...FixedAspect
    case 0: out[2] = out[3] = 1.f; break;                                   // InputResolution
    default: OPB_FAIL(OPB200_ERR_INVALID, "Unknown ScaleMode selected.");
    }
    return OPB200_OK;
}

extern "C" opb200_status opb200_scale_keypoints(float* keypoints, const int* count, int n_frames, int capacity,
                                                int parts, float scale_x, float scale_y, float offset_x,
                                                float offset_y, void* stream)
{
    using namespace opb;
    OPB_REQUIRE(keypoints, "scale_keypoints: null argument");
    OPB_REQUIRE(n_frames > 0 && n_frames <= 65535 && capacity > 0 && parts > 0, "scale_keypoints: empty people arrays");
    // keypointScaler.cpp:90-97 + the early-outs of both scaleKeypoints2d overloads
    const bool withOffset = !(offset_x == 0.f && offset_y == 0.f);
    if (!withOffset && scale_x == 1.f && scale_y == 1.f)
        return OPB200_OK;
    const dim3 grid((capacity * parts + kPeopleThreads - 1) / kPeopleThreads, n_frames);
    scaleKernel<<<grid, kPeopleThreads, 0, static_cast<cudaStream_t>(stream)>>>(keypoints, count, capacity, parts,
                                                                               scale_x, scale_y, offset_x, offset_y,
                                                                               withOffset ? 1 : 0);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}

extern "C" opb200_status opb200_hand_face_rectangles(float* hand_rectangles, float* face_rectangles,
                                                     const float* keypoints, const int* count, int n_frames,
                                                     int capacity, int pose_model, void* stream)
{
    using namespace opb;
    OPB_REQUIRE(keypoints && count && (hand_rectangles || face_rectangles), "hand_face_rectangles: null argument");
    OPB_REQUIRE(n_frames > 0 && n_frames <= 65535 && capacity > 0, "hand_face_rectangles: empty people arrays");
    RectParts idx;
    int parts;
    // poseBodyPartMapStringToKey over POSE_BODY_25_BODY_PARTS / POSE_COCO_BODY_PARTS (poseParameters.cpp:8-68)
    if (pose_model == OPB200_POSE_BODY_25)
    {
        parts = 25;
        idx = {7, 6, 5, 4, 3, 2, 1, 0, 18, 17, 16, 15};
    }
    else if (pose_model == OPB200_POSE_COCO_18)
    {
        parts = 18;
        idx = {7, 6, 5, 4, 3, 2, 1, 0, 17, 16, 15, 14};
    }
    else
        OPB_FAIL(OPB200_ERR_INVALID, "hand_face_rectangles: BODY_25 and COCO_18 only");
    const dim3 grid((capacity + kPeopleThreads - 1) / kPeopleThreads, n_frames);
    rectanglesKernel<<<grid, kPeopleThreads, 0, static_cast<cudaStream_t>(stream)>>>(hand_rectangles, face_rectangles,
                                                                                    keypoints, count, capacity, parts,
                                                                                    idx);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}
