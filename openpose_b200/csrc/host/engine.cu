// Multi-GPU body-pose engine behind the C ABI (opb200_engine_*): frames in, people out, frame order kept.
//
// Replaces, for the hot path, the reference's worker topology: one PoseExtractor per GPU, each initialised and used on
// its own thread (include/openpose/wrapper/wrapperAuxiliary.hpp:181-219, 326-338), fed round-robin from one queue and
// re-ordered by frame id afterwards (include/openpose/thread/wQueueOrderer.hpp:12-80). Frames are independent, so they
// are sharded by batch across the devices; there is no cross-GPU exchange and no collective on this path.
//
// Per device: one host thread, one compute stream, one copy stream, `lanes` batches in flight. A lane owns a pinned
// host staging buffer the producer writes u8 BGR frames into (opb200_engine_acquire: no extra host copy), its device
// copy, and pinned result buffers; heat-maps, peaks and kernel scratch are shared by the lanes of a device because
// compute is serialised on the compute stream. Per batch the worker enqueues, without synchronising:
//   copy stream:    H2D of the u8 frames
//   compute stream: opb200_preprocess (u8 HWC -> fp32 CHW net input, written straight into the net's input buffer)
//                   -> opb200_net_forward -> opb200_resize_merge -> opb200_nms -> opb200_connect -> D2H of
//                   keypoints / scores / counts
// and only then waits for the batch issued BEFORE it, so the H2D of batch i+1 overlaps the compute of batch i.
// opb200_engine_next hands batches back in submission (= frame id) order whatever device finished first.
#include "../common.cuh"
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace
{
    using namespace opb;

    enum LaneState
    {
        LANE_FREE,
        LANE_ACQUIRED,   // the producer is filling the staging buffer
        LANE_PENDING,    // committed, not yet enqueued on the device
        LANE_ISSUED,     // enqueued on the streams
        LANE_DONE        // results are in the pinned result buffers
    };

    struct Lane
    {
        unsigned char* hIn = nullptr;     // pinned: batch x frame_h x frame_w x 3
        unsigned char* dIn = nullptr;
        float *dKp = nullptr, *dSc = nullptr;
        int* dCnt = nullptr;
        float *hKp = nullptr, *hSc = nullptr;   // pinned
        int* hCnt = nullptr;
        cudaEvent_t h2d = nullptr, done = nullptr;
        int n = 0;
        unsigned long long firstId = 0;
        LaneState state = LANE_FREE;
    };

    struct Device
    {
        int gpu = 0;
        opb200_net* net = nullptr;
        cudaStream_t compute = nullptr, copy = nullptr;
        float *dHeat = nullptr, *dPeaks = nullptr;
        void *nmsScratch = nullptr, *connectScratch = nullptr, *maxScratch = nullptr;
        std::vector<Lane> lanes;
        std::deque<int> pending, issued;
        std::thread worker;
    };
}

struct opb200_engine
{
    opb200_engine_config cfg;
    bool cropMode = false;                    // hand / face crops of one frame per batch (opb200_crop_engine_create)
    opb200_crop_engine_config crop;
    int parts = 0, channels = 0;
    size_t frameBytes = 0, kpFloats = 0;      // bytes of one staged input unit; result floats per frame (per crop)
    size_t stagingBytes = 0, matOffset = 0;   // crop mode: frame | 2x3 double matrices of up to `batch` crops
    std::vector<std::unique_ptr<Device>> devices;
    std::mutex mtx;
    std::condition_variable cv;
    std::deque<std::pair<int, int>> order;   // (device, lane) in submission order
    int nextDevice = 0;
    int acquiredDevice = -1, acquiredLane = -1;
    bool stop = false;
    std::string error;                        // first failure of a worker thread
};

namespace
{
    // everything one batch needs, enqueued without a host synchronisation
    // crop mode: one u8 frame + the affine maps of ln.n crops -> opb200_crop_warp (into the net's input buffer) ->
    // opb200_net_forward -> opb200_resize_merge (x8) -> opb200_maximum -> D2H of (n, parts, 3) peaks in crop coordinates:
    // the device part of HandExtractorCaffe::forwardPass / FaceExtractorCaffe::forwardPass for all crops of the batch
    // at once (reference: src/openpose/hand/handExtractorCaffe.cpp:305-464, src/openpose/face/faceExtractorCaffe.cpp:175-300)
    opb200_status issueCropBatch(opb200_engine* e, Device& d, Lane& ln)
    {
        const opb200_crop_engine_config& c = e->crop;
        const int n = ln.n;
        OPB_CUDA(cudaMemcpyAsync(ln.dIn, ln.hIn, e->stagingBytes, cudaMemcpyHostToDevice, d.copy));
        OPB_CUDA(cudaEventRecord(ln.h2d, d.copy));
        OPB_CUDA(cudaStreamWaitEvent(d.compute, ln.h2d, 0));
        // the net always runs at the full batch shape (one CUDA graph, no re-planning when the number of crops changes from
        // frame to frame); slots beyond n keep whatever an earlier batch left there and their outputs are ignored
        const int inShape[4] = {c.batch, 3, c.net_side, c.net_side};
        float* netIn = nullptr;
        opb200_status s = opb200_net_input_buffer(d.net, inShape, &netIn, d.compute);
        if (s != OPB200_OK)
            return s;
        s = opb200_crop_warp(netIn, ln.dIn, reinterpret_cast<const double*>(ln.dIn + e->matOffset), n, c.frame_h, c.frame_w,
                             c.net_side, c.net_side, d.compute);
        if (s != OPB200_OK)
            return s;
        s = opb200_net_forward(d.net, netIn, 1, inShape, d.compute);
        if (s != OPB200_OK)
            return s;
        const float* netOut = nullptr;
        int outShape[4];
        s = opb200_net_output(d.net, &netOut, outShape);
        if (s != OPB200_OK)
            return s;
        OPB_REQUIRE(outShape[1] == c.heat_channels && outShape[2] * 8 == c.net_side && outShape[3] * 8 == c.net_side,
                    "crop engine: the net's output is not (heat_channels, net_side / 8, net_side / 8)");
        const int heatShape[4] = {n, c.heat_channels, c.net_side, c.net_side};
        const int peakShape[4] = {n, 1, c.heat_channels - 1, 3};
        const float* srcs[1] = {netOut};
        const int srcShapes[1][4] = {{n, outShape[1], outShape[2], outShape[3]}};
        const float ratios[1] = {1.f};
        s = opb200_resize_merge(d.dHeat, srcs, srcShapes, ratios, 1, heatShape, d.compute);
        if (s != OPB200_OK)
            return s;
        s = opb200_maximum(ln.dKp, d.maxScratch, d.dHeat, peakShape, heatShape, d.compute);
        if (s != OPB200_OK)
            return s;
        OPB_CUDA(cudaMemcpyAsync(ln.hKp, ln.dKp, sizeof(float) * e->kpFloats * n, cudaMemcpyDeviceToHost, d.compute));
        OPB_CUDA(cudaEventRecord(ln.done, d.compute));
        return OPB200_OK;
    }

    opb200_status issueBatch(opb200_engine* e, Device& d, Lane& ln)
    {
        if (e->cropMode)
            return issueCropBatch(e, d, ln);
        const opb200_engine_config& c = e->cfg;
        const int B = ln.n;
        OPB_CUDA(cudaMemcpyAsync(ln.dIn, ln.hIn, e->frameBytes * B, cudaMemcpyHostToDevice, d.copy));
        OPB_CUDA(cudaEventRecord(ln.h2d, d.copy));
        OPB_CUDA(cudaStreamWaitEvent(d.compute, ln.h2d, 0));
        const int inShape[4] = {B, 3, c.net_h, c.net_w};
        float* netIn = nullptr;
        opb200_status s = opb200_net_input_buffer(d.net, inShape, &netIn, d.compute);
        if (s != OPB200_OK)
            return s;
        s = opb200_preprocess(netIn, ln.dIn, B, c.frame_h, c.frame_w, c.net_h, c.net_w, c.scale_input_to_net, d.compute);
        if (s != OPB200_OK)
            return s;
        s = opb200_net_forward(d.net, netIn, 1, inShape, d.compute);
        if (s != OPB200_OK)
            return s;
        const float* netOut = nullptr;
        int outShape[4];
        s = opb200_net_output(d.net, &netOut, outShape);
        if (s != OPB200_OK)
            return s;
        const int heatShape[4] = {B, e->channels, c.net_h, c.net_w};
        const int peaksShape[4] = {B, e->parts, c.connect.max_peaks + 1, 3};
        const float* srcs[1] = {netOut};
        const int srcShapes[1][4] = {{outShape[0], outShape[1], outShape[2], outShape[3]}};
        const float ratios[1] = {1.f};
        s = opb200_resize_merge(d.dHeat, srcs, srcShapes, ratios, 1, heatShape, d.compute);
        if (s != OPB200_OK)
            return s;
        const float off = (float)(0.5 / (double)c.connect.scale_factor);   // poseExtractorCaffe.cpp:316
        s = opb200_nms(d.dPeaks, d.nmsScratch, d.dHeat, c.nms_threshold, off, off, peaksShape, heatShape,
                       c.connect.semantics, d.compute);
        if (s != OPB200_OK)
            return s;
        s = opb200_connect(ln.dKp, ln.dSc, ln.dCnt, c.capacity, d.connectScratch, d.dHeat, d.dPeaks, heatShape, &c.connect,
                           d.compute);
        if (s != OPB200_OK)
            return s;
        OPB_CUDA(cudaMemcpyAsync(ln.hKp, ln.dKp, sizeof(float) * e->kpFloats * B, cudaMemcpyDeviceToHost, d.compute));
        OPB_CUDA(cudaMemcpyAsync(ln.hSc, ln.dSc, sizeof(float) * c.capacity * B, cudaMemcpyDeviceToHost, d.compute));
        OPB_CUDA(cudaMemcpyAsync(ln.hCnt, ln.dCnt, sizeof(int) * B, cudaMemcpyDeviceToHost, d.compute));
        OPB_CUDA(cudaEventRecord(ln.done, d.compute));
        return OPB200_OK;
    }

    void fail(opb200_engine* e, const char* what)
    {
        std::lock_guard<std::mutex> lock(e->mtx);
        if (e->error.empty())
            e->error = what;
        e->stop = true;
        e->cv.notify_all();
    }

    void workerLoop(opb200_engine* e, Device* d)
    {
        if (cudaSetDevice(d->gpu) != cudaSuccess)
            return fail(e, "engine worker: cudaSetDevice failed");
        for (;;)
        {
            int toIssue = -1;
            {
                std::unique_lock<std::mutex> lock(e->mtx);
                auto ready = [&] { return e->stop || !d->pending.empty(); };
                if (d->issued.empty())
                    e->cv.wait(lock, ready);
                else   // batches in flight: keep polling their events while waiting for the next one to enqueue
                    e->cv.wait_for(lock, std::chrono::microseconds(50), ready);
                if (e->stop && d->pending.empty() && d->issued.empty())
                    return;
                if (!d->pending.empty())
                {
                    toIssue = d->pending.front();
                    d->pending.pop_front();
                }
            }
            if (toIssue >= 0)
            {
                if (issueBatch(e, *d, d->lanes[toIssue]) != OPB200_OK)
                    return fail(e, opb200_last_error());
                std::lock_guard<std::mutex> lock(e->mtx);
                d->lanes[toIssue].state = LANE_ISSUED;
                d->issued.push_back(toIssue);
            }
            // retire finished batches in order; never block here (the next batch must be enqueued while this one runs)
            for (;;)
            {
                int head = -1;
                {
                    std::lock_guard<std::mutex> lock(e->mtx);
                    if (!d->issued.empty())
                        head = d->issued.front();
                }
                if (head < 0)
                    break;
                const cudaError_t q = cudaEventQuery(d->lanes[head].done);
                if (q == cudaErrorNotReady)
                    break;
                if (q != cudaSuccess)
                    return fail(e, "engine worker: a batch failed on the device");
                std::lock_guard<std::mutex> lock(e->mtx);
                d->lanes[head].state = LANE_DONE;
                d->issued.pop_front();
                e->cv.notify_all();
            }
        }
    }

    void destroyDevice(Device& d)
    {
        cudaSetDevice(d.gpu);
        if (d.compute)
            cudaStreamSynchronize(d.compute);
        if (d.copy)
            cudaStreamSynchronize(d.copy);
        for (auto& ln : d.lanes)
        {
            cudaFreeHost(ln.hIn);
            cudaFree(ln.dIn);
            cudaFree(ln.dKp);
            cudaFree(ln.dSc);
            cudaFree(ln.dCnt);
            cudaFreeHost(ln.hKp);
            cudaFreeHost(ln.hSc);
            cudaFreeHost(ln.hCnt);
            if (ln.h2d)
                cudaEventDestroy(ln.h2d);
            if (ln.done)
                cudaEventDestroy(ln.done);
        }
        cudaFree(d.dHeat);
        cudaFree(d.dPeaks);
        cudaFree(d.nmsScratch);
        cudaFree(d.connectScratch);
        cudaFree(d.maxScratch);
        if (d.net)
            opb200_net_destroy(d.net);
        if (d.compute)
            cudaStreamDestroy(d.compute);
        if (d.copy)
            cudaStreamDestroy(d.copy);
    }

    opb200_status createDevice(opb200_engine* e, Device& d, const char* prototxt, const char* caffemodel)
    {
        const opb200_engine_config& c = e->cfg;
        OPB_CUDA(cudaSetDevice(d.gpu));
        opb200_status s = opb200_net_create(&d.net, prototxt, caffemodel, d.gpu, "net_output");
        if (s != OPB200_OK)
            return s;
        OPB_CUDA(cudaStreamCreateWithFlags(&d.compute, cudaStreamNonBlocking));
        OPB_CUDA(cudaStreamCreateWithFlags(&d.copy, cudaStreamNonBlocking));
        if (e->cropMode)
        {
            const opb200_crop_engine_config& cc = e->crop;
            const size_t n = (size_t)cc.batch;
            const int heatShape[4] = {cc.batch, cc.heat_channels, cc.net_side, cc.net_side};
            const int peakShape[4] = {cc.batch, 1, cc.heat_channels - 1, 3};
            OPB_CUDA(cudaMalloc(&d.dHeat, sizeof(float) * n * cc.heat_channels * cc.net_side * cc.net_side));
            OPB_CUDA(cudaMalloc(&d.maxScratch, opb200_maximum_scratch_bytes(peakShape, heatShape)));
            {
                const int inShape[4] = {cc.batch, 3, cc.net_side, cc.net_side};
                float* netIn = nullptr;
                const opb200_status si = opb200_net_input_buffer(d.net, inShape, &netIn, d.compute);
                if (si != OPB200_OK)
                    return si;
                OPB_CUDA(cudaMemsetAsync(netIn, 0, sizeof(float) * n * 3 * cc.net_side * cc.net_side, d.compute));
                OPB_CUDA(cudaStreamSynchronize(d.compute));
            }
            d.lanes.resize(cc.lanes);
            for (auto& ln : d.lanes)
            {
                OPB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ln.hIn), e->stagingBytes, cudaHostAllocDefault));
                OPB_CUDA(cudaMalloc(&ln.dIn, e->stagingBytes));
                OPB_CUDA(cudaMalloc(&ln.dKp, sizeof(float) * e->kpFloats * n));
                OPB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ln.hKp), sizeof(float) * e->kpFloats * n, cudaHostAllocDefault));
                OPB_CUDA(cudaEventCreateWithFlags(&ln.h2d, cudaEventDisableTiming));
                OPB_CUDA(cudaEventCreateWithFlags(&ln.done, cudaEventDisableTiming));
            }
            return OPB200_OK;
        }
        const size_t B = (size_t)c.batch;
        const int heatShape[4] = {c.batch, e->channels, c.net_h, c.net_w};
        const int peaksShape[4] = {c.batch, e->parts, c.connect.max_peaks + 1, 3};
        OPB_CUDA(cudaMalloc(&d.dHeat, sizeof(float) * B * e->channels * c.net_h * c.net_w));
        OPB_CUDA(cudaMalloc(&d.dPeaks, sizeof(float) * B * e->parts * (c.connect.max_peaks + 1) * 3));
        OPB_CUDA(cudaMemset(d.dPeaks, 0, sizeof(float) * B * e->parts * (c.connect.max_peaks + 1) * 3));
        OPB_CUDA(cudaMalloc(&d.nmsScratch, opb200_nms_scratch_bytes(peaksShape, heatShape)));
        OPB_CUDA(cudaMalloc(&d.connectScratch, opb200_connect_scratch_bytes(&c.connect, c.batch)));
        d.lanes.resize(c.lanes);
        for (auto& ln : d.lanes)
        {
            OPB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ln.hIn), e->frameBytes * B, cudaHostAllocDefault));
            OPB_CUDA(cudaMalloc(&ln.dIn, e->frameBytes * B));
            OPB_CUDA(cudaMalloc(&ln.dKp, sizeof(float) * e->kpFloats * B));
            OPB_CUDA(cudaMalloc(&ln.dSc, sizeof(float) * c.capacity * B));
            OPB_CUDA(cudaMalloc(&ln.dCnt, sizeof(int) * B));
            OPB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ln.hKp), sizeof(float) * e->kpFloats * B, cudaHostAllocDefault));
            OPB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ln.hSc), sizeof(float) * c.capacity * B, cudaHostAllocDefault));
            OPB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ln.hCnt), sizeof(int) * B, cudaHostAllocDefault));
            OPB_CUDA(cudaEventCreateWithFlags(&ln.h2d, cudaEventDisableTiming));
            OPB_CUDA(cudaEventCreateWithFlags(&ln.done, cudaEventDisableTiming));
        }
        return OPB200_OK;
    }
}

extern "C" opb200_status opb200_engine_create(opb200_engine** out, const char* prototxt, const char* caffemodel,
                                              const int* gpu_ids, int n_gpus, const opb200_engine_config* cfg)
{
    OPB_REQUIRE(out && prototxt && caffemodel && gpu_ids && cfg, "engine_create: null argument");
    *out = nullptr;
    OPB_REQUIRE(n_gpus >= 1 && n_gpus <= 64, "engine_create: between 1 and 64 devices");
    OPB_REQUIRE(cfg->pose_model == OPB200_POSE_BODY_25 || cfg->pose_model == OPB200_POSE_COCO_18,
                "engine_create: BODY_25 and COCO_18 only");
    OPB_REQUIRE(cfg->batch >= 1 && cfg->lanes >= 1 && cfg->lanes <= 8 && cfg->capacity >= 1, "engine_create: bad batch / lanes / capacity");
    OPB_REQUIRE(cfg->net_w > 0 && cfg->net_h > 0 && cfg->net_w % 8 == 0 && cfg->net_h % 8 == 0,
                "engine_create: the net resolution must be a positive multiple of 8");
    OPB_REQUIRE(cfg->frame_w > 0 && cfg->frame_h > 0 && cfg->scale_input_to_net > 0.f && cfg->connect.scale_factor > 0.f,
                "engine_create: bad frame size or scale");
    OPB_REQUIRE(cfg->connect.pose_model == cfg->pose_model && cfg->connect.max_peaks >= 1, "engine_create: connect params inconsistent");
    std::unique_ptr<opb200_engine> e(new opb200_engine());
    e->cfg = *cfg;
    e->parts = cfg->pose_model == OPB200_POSE_BODY_25 ? 25 : 18;
    e->channels = cfg->pose_model == OPB200_POSE_BODY_25 ? 78 : 57;
    e->frameBytes = (size_t)cfg->frame_h * cfg->frame_w * 3;
    e->kpFloats = (size_t)cfg->capacity * e->parts * 3;
    for (int i = 0; i < n_gpus; ++i)
    {
        e->devices.emplace_back(new Device());
        e->devices.back()->gpu = gpu_ids[i];
        const opb200_status s = createDevice(e.get(), *e->devices.back(), prototxt, caffemodel);
        if (s != OPB200_OK)
        {
            for (auto& d : e->devices)
                destroyDevice(*d);
            return s;
        }
    }
    for (auto& d : e->devices)
        d->worker = std::thread(workerLoop, e.get(), d.get());
    *out = e.release();
    return OPB200_OK;
}

extern "C" opb200_status opb200_crop_engine_create(opb200_engine** out, const char* prototxt, const char* caffemodel,
                                                   const int* gpu_ids, int n_gpus, const opb200_crop_engine_config* cfg)
{
    OPB_REQUIRE(out && prototxt && caffemodel && gpu_ids && cfg, "crop_engine_create: null argument");
    *out = nullptr;
    OPB_REQUIRE(n_gpus >= 1 && n_gpus <= 64, "crop_engine_create: between 1 and 64 devices");
    OPB_REQUIRE(cfg->batch >= 1 && cfg->lanes >= 1 && cfg->lanes <= 8, "crop_engine_create: bad batch / lanes");
    OPB_REQUIRE(cfg->heat_channels >= 2, "crop_engine_create: heat_channels counts the background channel too (22 hand, 71 face)");
    OPB_REQUIRE(cfg->net_side > 0 && cfg->net_side % 8 == 0, "crop_engine_create: the net side must be a positive multiple of 8");
    OPB_REQUIRE(cfg->frame_w > 0 && cfg->frame_h > 0, "crop_engine_create: bad frame size");
    std::unique_ptr<opb200_engine> e(new opb200_engine());
    e->cropMode = true;
    e->crop = *cfg;
    e->cfg = opb200_engine_config();
    e->cfg.batch = cfg->batch;
    e->cfg.lanes = cfg->lanes;
    e->parts = cfg->heat_channels - 1;
    e->channels = cfg->heat_channels;
    e->frameBytes = (size_t)cfg->frame_h * cfg->frame_w * 3;
    e->matOffset = (e->frameBytes + 15) & ~(size_t)15;
    e->stagingBytes = e->matOffset + sizeof(double) * 6 * (size_t)cfg->batch;
    e->kpFloats = (size_t)e->parts * 3;
    for (int i = 0; i < n_gpus; ++i)
    {
        e->devices.emplace_back(new Device());
        e->devices.back()->gpu = gpu_ids[i];
        const opb200_status s = createDevice(e.get(), *e->devices.back(), prototxt, caffemodel);
        if (s != OPB200_OK)
        {
            for (auto& d : e->devices)
                destroyDevice(*d);
            return s;
        }
    }
    for (auto& d : e->devices)
        d->worker = std::thread(workerLoop, e.get(), d.get());
    *out = e.release();
    return OPB200_OK;
}

extern "C" size_t opb200_crop_engine_matrix_offset(opb200_engine* e) { return (e && e->cropMode) ? e->matOffset : 0; }

extern "C" void opb200_engine_destroy(opb200_engine* e)
{
    if (!e)
        return;
    {
        std::lock_guard<std::mutex> lock(e->mtx);
        e->stop = true;
        e->cv.notify_all();
    }
    for (auto& d : e->devices)
        if (d->worker.joinable())
            d->worker.join();
    for (auto& d : e->devices)
        destroyDevice(*d);
    delete e;
}

extern "C" opb200_status opb200_engine_acquire(opb200_engine* e, unsigned char** staging)
{
    OPB_REQUIRE(e && staging, "engine_acquire: null argument");
    std::unique_lock<std::mutex> lock(e->mtx);
    OPB_REQUIRE(e->acquiredLane < 0, "engine_acquire: the previous buffer was not committed");
    const int dev = e->nextDevice;
    Device& d = *e->devices[dev];
    int lane = -1;
    e->cv.wait(lock, [&] {
        if (e->stop)
            return true;
        for (size_t i = 0; i < d.lanes.size(); ++i)
            if (d.lanes[i].state == LANE_FREE)
            {
                lane = (int)i;
                return true;
            }
        return false;
    });
    if (lane < 0)
    {
        setError("engine: %s", e->error.empty() ? "stopped" : e->error.c_str());
        return OPB200_ERR_CUDA;
    }
    d.lanes[lane].state = LANE_ACQUIRED;
    e->acquiredDevice = dev;
    e->acquiredLane = lane;
    *staging = d.lanes[lane].hIn;
    return OPB200_OK;
}

extern "C" opb200_status opb200_engine_commit(opb200_engine* e, unsigned long long first_frame_id, int n_frames)
{
    OPB_REQUIRE(e != nullptr, "engine_commit: null engine");
    std::lock_guard<std::mutex> lock(e->mtx);
    OPB_REQUIRE(e->acquiredLane >= 0, "engine_commit: no acquired buffer");
    OPB_REQUIRE(n_frames >= 1 && n_frames <= e->cfg.batch, "engine_commit: between 1 and `batch` frames");
    Device& d = *e->devices[e->acquiredDevice];
    Lane& ln = d.lanes[e->acquiredLane];
    ln.n = n_frames;
    ln.firstId = first_frame_id;
    ln.state = LANE_PENDING;
    d.pending.push_back(e->acquiredLane);
    e->order.emplace_back(e->acquiredDevice, e->acquiredLane);
    e->acquiredDevice = e->acquiredLane = -1;
    e->nextDevice = (e->nextDevice + 1) % (int)e->devices.size();   // batches go round the devices
    e->cv.notify_all();
    return OPB200_OK;
}

extern "C" opb200_status opb200_engine_submit(opb200_engine* e, unsigned long long first_frame_id,
                                              const unsigned char* frames, int n_frames)
{
    OPB_REQUIRE(e && frames, "engine_submit: null argument");
    OPB_REQUIRE(!e->cropMode, "engine_submit: a crop engine takes opb200_crop_engine_submit");
    OPB_REQUIRE(n_frames >= 1 && n_frames <= e->cfg.batch, "engine_submit: between 1 and `batch` frames");
    unsigned char* staging = nullptr;
    const opb200_status s = opb200_engine_acquire(e, &staging);
    if (s != OPB200_OK)
        return s;
    std::memcpy(staging, frames, e->frameBytes * (size_t)n_frames);
    return opb200_engine_commit(e, first_frame_id, n_frames);
}

extern "C" int opb200_engine_in_flight(opb200_engine* e)
{
    if (!e)
        return 0;
    std::lock_guard<std::mutex> lock(e->mtx);
    return (int)e->order.size();
}

extern "C" opb200_status opb200_engine_next(opb200_engine* e, unsigned long long* first_frame_id, int* n_frames,
                                            float* keypoints, float* scores, int* counts)
{
    OPB_REQUIRE(e && first_frame_id && n_frames && keypoints && scores && counts, "engine_next: null argument");
    OPB_REQUIRE(!e->cropMode, "engine_next: a crop engine takes opb200_crop_engine_next");
    std::unique_lock<std::mutex> lock(e->mtx);
    OPB_REQUIRE(!e->order.empty(), "engine_next: nothing was submitted");
    const std::pair<int, int> head = e->order.front();
    Lane& ln = e->devices[head.first]->lanes[head.second];
    e->cv.wait(lock, [&] { return e->stop || ln.state == LANE_DONE; });
    if (ln.state != LANE_DONE)
    {
        setError("engine: %s", e->error.empty() ? "stopped" : e->error.c_str());
        return OPB200_ERR_CUDA;
    }
    const int B = ln.n;
    *first_frame_id = ln.firstId;
    *n_frames = B;
    lock.unlock();   // the lane stays DONE (nobody else touches it) while its results are copied out
    std::memcpy(keypoints, ln.hKp, sizeof(float) * e->kpFloats * B);
    std::memcpy(scores, ln.hSc, sizeof(float) * e->cfg.capacity * B);
    std::memcpy(counts, ln.hCnt, sizeof(int) * B);
    lock.lock();
    ln.state = LANE_FREE;
    e->order.pop_front();
    e->cv.notify_all();
    return OPB200_OK;
}

extern "C" opb200_status opb200_crop_engine_submit(opb200_engine* e, unsigned long long frame_id, const unsigned char* frame,
                                                   const double* matrices, int n_crops)
{
    OPB_REQUIRE(e && frame && matrices, "crop_engine_submit: null argument");
    OPB_REQUIRE(e->cropMode, "crop_engine_submit: not a crop engine");
    OPB_REQUIRE(n_crops >= 1 && n_crops <= e->crop.batch, "crop_engine_submit: between 1 and `batch` crops");
    unsigned char* staging = nullptr;
    const opb200_status s = opb200_engine_acquire(e, &staging);
    if (s != OPB200_OK)
        return s;
    std::memcpy(staging, frame, e->frameBytes);
    std::memcpy(staging + e->matOffset, matrices, sizeof(double) * 6 * (size_t)n_crops);
    return opb200_engine_commit(e, frame_id, n_crops);
}

extern "C" opb200_status opb200_crop_engine_next(opb200_engine* e, unsigned long long* frame_id, int* n_crops, float* peaks)
{
    OPB_REQUIRE(e && frame_id && n_crops && peaks, "crop_engine_next: null argument");
    OPB_REQUIRE(e->cropMode, "crop_engine_next: not a crop engine");
    std::unique_lock<std::mutex> lock(e->mtx);
    OPB_REQUIRE(!e->order.empty(), "crop_engine_next: nothing was submitted");
    const std::pair<int, int> head = e->order.front();
    Lane& ln = e->devices[head.first]->lanes[head.second];
    e->cv.wait(lock, [&] { return e->stop || ln.state == LANE_DONE; });
    if (ln.state != LANE_DONE)
    {
        setError("engine: %s", e->error.empty() ? "stopped" : e->error.c_str());
        return OPB200_ERR_CUDA;
    }
    *frame_id = ln.firstId;
    *n_crops = ln.n;
    lock.unlock();
    std::memcpy(peaks, ln.hKp, sizeof(float) * e->kpFloats * (size_t)ln.n);
    lock.lock();
    ln.state = LANE_FREE;
    e->order.pop_front();
    e->cv.notify_all();
    return OPB200_OK;
}
