/* opb200.h -- C ABI of the B200-native OpenPose hot path (libopb200.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types, no exceptions. Every entry
 * point returns 0 on success or a non-zero opb200_status; opb200_last_error() then holds a message that
 * the C++ shim (openpose_b200/csrc/op/) turns into op::error(...) exactly like the reference's
 * try/catch -> error(e.what(), __LINE__, __FUNCTION__, __FILE__) convention
 * (reference: src/openpose/utilities/errorAndLog.cpp:158-233).
 *
 * "reference" paths below are relative to the zengjianping/openpose tree.
 * All data pointers are DEVICE pointers unless the name says host; `stream` is a cudaStream_t passed as
 * void* (NULL = legacy default stream). Nothing in here synchronises the device unless stated.
 */
#ifndef OPB200_H
#define OPB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int opb200_status;
#define OPB200_OK 0
#define OPB200_ERR_INVALID 1   /* bad argument (shape, threshold, too many scales ...) */
#define OPB200_ERR_CUDA 2      /* a CUDA runtime call or kernel launch failed */
#define OPB200_ERR_UNSUPPORTED 3
#define OPB200_ERR_IO 4        /* prototxt / caffemodel missing or malformed */

/* op::PoseModel values (reference: include/openpose/pose/enumClasses.hpp:9-30). */
#define OPB200_POSE_BODY_25 0
#define OPB200_POSE_COCO_18 1

/* Which of the reference's two divergent semantics to reproduce. */
#define OPB200_SEMANTICS_CPU 0  /* CPU_ONLY build: nmsBase.cpp, bodyPartConnectorBase.cpp createPeopleVector  */
#define OPB200_SEMANTICS_CUDA 1 /* CUDA build: nmsBase.cu borders, process() scores, pafVectorIntoPeopleVector */

#define OPB200_MAX_SCALES 8     /* reference: src/openpose/net/resizeAndMergeBase.cu:457-459 */

const char* opb200_last_error(void);   /* thread-local, valid until the next failing call on this thread */
int opb200_version(void);              /* ABI version, currently 1 */
int opb200_device_count(void);         /* replaces op::getCudaGpuNumber, reference: src/openpose/gpu/cuda.cpp:41-61 */

/* ---------------------------------------------------------------------------------------------------
 * resizeAndMerge -- replaces op::resizeAndMergeGpu<float>
 * (reference: include/openpose/net/resizeAndMergeBase.hpp:14-17, src/openpose/net/resizeAndMergeBase.cu:275-528).
 * dst (dst_shape = N,C,H,W) = average over scales of the bicubic (Catmull-Rom, cuda.hu:91-144) upsampling of
 * srcs[i] (src_shapes[i] = N_i,C,h_i,w_i). One scale: ratio 1 copies, otherwise the x/y ratios must be equal
 * integers (the reference only implements x8; any integer factor is accepted here). Several scales (<= 8):
 * per-scale factors derived from scale_ratios[i] (scaleInputToNetInputs) as resizeAndMergeBase.cu:379-380,475-477.
 * N: every frame n of dst is built from frame n of each source (the reference merges only N == 1).
 * `srcs` is a HOST array of device pointers.
 * ------------------------------------------------------------------------------------------------- */
opb200_status opb200_resize_merge(float* dst, const float* const* srcs, const int (*src_shapes)[4],
                                  const float* scale_ratios, int n_scales, const int dst_shape[4],
                                  void* stream);

/* ---------------------------------------------------------------------------------------------------
 * nms -- replaces op::nmsGpu<float> / op::nmsCpu<float>
 * (reference: include/openpose/net/nmsBase.hpp:8-17, src/openpose/net/nmsBase.cpp:110-170, nmsBase.cu:249-351).
 * heat (heat_shape = N,Ch,H,W): only channels [0, peaks_shape[1]) of each frame are scanned.
 * peaks (peaks_shape = N,parts,maxPeaks+1,3): row 0 col 0 = count, rows 1..count = (x, y, score); rows beyond
 * the count are left untouched, as in the reference.
 * scratch: device buffer of at least opb200_nms_scratch_bytes(...) bytes (the reference's int "kernel" blob,
 * nmsCaffe.cpp:123, is always large enough). mode = OPB200_SEMANTICS_*.
 * ------------------------------------------------------------------------------------------------- */
size_t opb200_nms_scratch_bytes(const int peaks_shape[4], const int heat_shape[4]);
opb200_status opb200_nms(float* peaks, void* scratch, const float* heat, float threshold, float offset_x,
                         float offset_y, const int peaks_shape[4], const int heat_shape[4], int mode,
                         void* stream);

/* opb200_nms_static -- opb200_nms without a scratch argument, for callers that cannot allocate (CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc). It has NO counterpart in the reference.
 * The scratch is a file-scope `static __device__` array of
 * OPB200_NMS_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until exit); the
 * call performs no cudaMalloc / cudaFree / allocating library call. Shapes whose opb200_nms_scratch_bytes(...)
 * exceeds the array are rejected with OPB200_ERR_INVALID (4 MiB = 5 BODY_25 or 7 COCO frames of 127 peaks).
 * All calls share the one array per device: what must start at zero is cleared on `stream` by every call, but two
 * calls that overlap on the GPU (different streams, no ordering between them) RACE -- serialise them.
 * Same arguments, results and errors as opb200_nms otherwise. */
#define OPB200_NMS_STATIC_SCRATCH_BYTES (4u << 20)
opb200_status opb200_nms_static(float* peaks, const float* heat, float threshold, float offset_x, float offset_y,
                                const int peaks_shape[4], const int heat_shape[4], int mode, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * maximum -- replaces op::maximumGpu<float> / op::maximumCpu<float>
 * (reference: include/openpose/net/maximumBase.hpp, src/openpose/net/maximumBase.cpp:8-57, maximumBase.cu:63-120).
 * source (N,C,H,W) -> target (N,1,C-1,3) = (x, y, value) of the first maximum in raster order of channels
 * 0..C-2 (the background channel is skipped). scratch: >= opb200_maximum_scratch_bytes(...) bytes.
 * ------------------------------------------------------------------------------------------------- */
size_t opb200_maximum_scratch_bytes(const int target_shape[4], const int source_shape[4]);
opb200_status opb200_maximum(float* target, void* scratch, const float* source, const int target_shape[4],
                             const int source_shape[4], void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Frame pre-processing -- replaces the device branch of op::CvMatToOpInput::createArray
 * (reference: src/openpose/core/cvMatToOpInput.cpp:108-148): op::reorderAndNormalize
 * (include/openpose/gpu/cuda.hpp, src/openpose/gpu/cuda.cu:11-25,69-83) followed by op::resizeAndPadRbgGpu
 * (include/openpose/net/resizeAndMergeBase.hpp, src/openpose/net/resizeAndMergeBase.cu:56-78,531-551), as one
 * kernel without the float scratch frame. src: n_frames BGR frames (N,Hs,Ws,3) u8 on the device;
 * dst: (N,3,Ht,Wt) fp32 = bicubic sample of (v/256 - 0.5) at ((x+0.5)/scale-0.5, (y+0.5)/scale-0.5) inside
 * [0,Ws*scale) x [0,Hs*scale), 0 elsewhere -- the net input op::Net::forwardPass / opb200_net_forward takes.
 * ------------------------------------------------------------------------------------------------- */
opb200_status opb200_preprocess(float* dst, const unsigned char* src, int n_frames, int src_h, int src_w,
                                int dst_h, int dst_w, float scale, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Hand / face crops -- replaces cv::warpAffine(INTER_LINEAR | WARP_INVERSE_MAP, BORDER_CONSTANT 0) +
 * op::uCharCvMatToFloatPtr(normalize = 1) in HandExtractorCaffe cropFrame (reference:
 * src/openpose/hand/handExtractorCaffe.cpp:44-76) and FaceExtractorCaffe::forwardPass
 * (src/openpose/face/faceExtractorCaffe.cpp:133-149; src/openpose/utilities/openCv.cpp:57-130).
 * frame: one BGR frame (Hs,Ws,3) u8 on the device; matrices: n_crops 2x3 row-major double inverse maps (crop pixel
 * -> frame position; the cv::Mat the reference builds) on the device; dst (n_crops,3,Ht,Wt) fp32 = v/256 - 0.5 of
 * OpenCV's 8-bit fixed-point bilinear sample. All crops of a frame in one launch.
 * ------------------------------------------------------------------------------------------------- */
opb200_status opb200_crop_warp(float* dst, const unsigned char* frame, const double* matrices, int n_crops,
                               int src_h, int src_w, int dst_h, int dst_w, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Heat-map export -- replaces the per-group cudaMemcpy + host remap loops of op::PoseExtractorNet::getHeatMapsCopy
 * (reference: src/openpose/pose/poseExtractorNet.cpp:106-243; the hand / face extractors apply the body-part rules,
 * handExtractorCaffe.cpp / faceExtractorCaffe.cpp updateHeatMapsForPerson). heat (N,C,H,W): channels
 * [0,parts) body parts, then 1 background channel when has_bkg, then paf_channels PAF channels. types_mask selects
 * the groups (OPB200_HEATMAP_*, bit = 1 << op::HeatMapType); scale_mode is the op::ScaleMode value (3..8:
 * ZeroToOne, ZeroToOneFixedAspect, PlusMinusOne, PlusMinusOneFixedAspect, UnsignedChar, NoScale; 0..2 clamp only,
 * the reference's final else branch).
 * dst (N, opb200_heatmaps_export_channels(...), H, W) fp32: a device buffer, or pinned host memory mapped into the
 * device address space (cudaHostAlloc / cudaHostRegister), which the kernel writes directly.
 * ------------------------------------------------------------------------------------------------- */
#define OPB200_HEATMAP_PARTS 1
#define OPB200_HEATMAP_BACKGROUND 2
#define OPB200_HEATMAP_PAFS 4
int opb200_heatmaps_export_channels(int parts, int has_bkg, int paf_channels, int types_mask);
opb200_status opb200_heatmaps_export(float* dst, const float* heat, const int heat_shape[4], int parts, int has_bkg,
                                     int paf_channels, int types_mask, int scale_mode, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * bodyPartConnector -- replaces op::connectBodyPartsGpu<float> / op::connectBodyPartsCpu<float>
 * (reference: include/openpose/net/bodyPartConnectorBase.hpp:8-24, src/openpose/net/bodyPartConnectorBase.cpp:
 * 12-75 getScoreAB, 157-472 createPeopleVector, 721-884 filter, 887-934 output; .cu:15-66,108-145 scores).
 * Everything (PAF line integrals, per-limb sort + greedy matching, person assembly, filtering, output) runs
 * on the device; the host only reads the final arrays.
 *   heat   (N,Ch,H,W) device: full heat-map blob (PAF channels at parts+1+mapIdx)
 *   peaks  (N,parts,max_peaks+1,3) device: output of opb200_nms
 *   out_keypoints (N,capacity,parts,3), out_scores (N,capacity), out_count (N) int32: device
 * out_count[n] = number of people found (may exceed capacity; only the first `capacity` are written).
 * ------------------------------------------------------------------------------------------------- */
typedef struct opb200_connect_params
{
    int pose_model;               /* OPB200_POSE_* */
    int max_peaks;                /* POSE_MAX_PEOPLE = 127 */
    float inter_min_above_threshold;
    float inter_threshold;
    int min_subset_cnt;
    float min_subset_score;
    float default_nms_threshold;
    float scale_factor;           /* scaleNetToOutput */
    int maximize_positives;
    int semantics;                /* OPB200_SEMANTICS_* */
} opb200_connect_params;

size_t opb200_connect_scratch_bytes(const opb200_connect_params* p, int n_frames);
opb200_status opb200_connect(float* out_keypoints, float* out_scores, int* out_count, int capacity,
                             void* scratch, const float* heat, const float* peaks, const int heat_shape[4],
                             const opb200_connect_params* p, void* stream);
/* Score cube only (pairs, max_peaks, max_peaks) per frame, as pafScoreKernel writes it
 * (reference: src/openpose/net/bodyPartConnectorBase.cu:108-145); for tests and the CUDA-semantics path. */
opb200_status opb200_paf_scores(float* cube, const float* heat, const float* peaks, const int heat_shape[4],
                                const opb200_connect_params* p, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * People post-steps on the connector's output layout (keypoints (N,capacity,parts,3), scores (N,capacity),
 * count (N) int32, all on the device), so body -> rectangles -> crops -> hand / face stays on the device.
 *
 * opb200_keep_top_people -- replaces op::KeepTopNPeople::keepTopPeople
 *   (reference: include/openpose/core/keepTopNPeople.hpp, src/openpose/core/keepTopNPeople.cpp:17-91):
 *   score * sqrt(area of the box around parts with score > 0.05), the number_people_max best kept in their
 *   original order, ties on the threshold filled in order; frames with <= number_people_max people (or
 *   number_people_max <= 0) are copied unchanged. out_* have the input layout; out_scores carries the kept
 *   people's scores (the reference leaves the score array untouched).
 * opb200_keypoint_scale_and_offset / opb200_scale_keypoints -- replace op::KeypointScaler::scale
 *   (reference: src/openpose/core/keypointScaler.cpp:7-104; op::scaleKeypoints2d,
 *   src/openpose/utilities/keypoint.cpp:104-168): host helper -> {offset x, offset y, scale x, scale y} for an
 *   op::ScaleMode, then x*sx[+ox], y*sy[+oy] in place (count may be NULL = all `capacity` rows, e.g. for hand /
 *   face keypoint arrays).
 * opb200_hand_face_rectangles -- replaces op::HandDetector::detectHands and op::FaceDetector::detectFaces
 *   (reference: src/openpose/hand/handDetector.cpp:9-46,136-164, src/openpose/face/faceDetector.cpp:22-141):
 *   hand_rectangles (N,capacity,2,4) = left, right (x, y, w, h); face_rectangles (N,capacity,4); either may be NULL.
 * ------------------------------------------------------------------------------------------------- */
opb200_status opb200_keep_top_people(float* out_keypoints, float* out_scores, int* out_count,
                                     const float* keypoints, const float* scores, const int* count, int n_frames,
                                     int capacity, int parts, int number_people_max, void* stream);
opb200_status opb200_keypoint_scale_and_offset(float out[4], int scale_mode, double scale_input_to_output,
                                               double scale_net_to_output, int producer_w, int producer_h);
opb200_status opb200_scale_keypoints(float* keypoints, const int* count, int n_frames, int capacity, int parts,
                                     float scale_x, float scale_y, float offset_x, float offset_y, void* stream);
opb200_status opb200_hand_face_rectangles(float* hand_rectangles, float* face_rectangles, const float* keypoints,
                                          const int* count, int n_frames, int capacity, int pose_model,
                                          void* stream);

/* ---------------------------------------------------------------------------------------------------
 * NetB200 -- replaces op::NetCaffe behind op::Net
 * (reference: include/openpose/net/net.hpp:8-18, include/openpose/net/netCaffe.hpp:12-13,
 *  src/openpose/net/netCaffe.cpp:45-99 constructor checks, :154-210 initializationOnThread,
 *  :212-261 forwardPass, :263-278 getOutputBlobArray).
 * opb200_net_create = constructor + initializationOnThread: checks both files exist (same messages), parses
 * the prototxt and the caffemodel (weights matched by layer name), plans the graph and uploads bf16 weights
 * to device gpu_id. last_blob_name NULL/"" = "net_output"; an unknown name fails with NetCaffe's message.
 * opb200_net_forward = forwardPass: input (N,3,H,W) fp32 NCHW on the host (input_on_device = 0; copied
 * inside the call, stream-ordered) or on the device; re-plans buffers when the shape changes; enqueues
 * the forward pass stream-ordered after `stream` and makes `stream` wait for it (no host synchronisation).
 * opb200_net_output = getOutputBlobArray: device pointer + shape of the fp32 NCHW output blob, owned by the
 * net, valid until the next reshaping forward or destroy.
 * Convolutions run in bf16 with fp32 accumulation on tcgen05 tensor cores; supported layer types:
 * Convolution (stride 1, "same" odd kernels), ReLU, PReLU, MAX Pooling (no pad), Concat (axis 1), Input.
 * ------------------------------------------------------------------------------------------------- */
typedef struct opb200_net opb200_net;
typedef struct opb200_net_info_t
{
    int conv_layers;
    int kernel_launches_per_forward;
    double flops_per_forward;     /* algorithmic conv FLOPs at the current shape: 2*Cout*Cin*k*k*H*W*N */
    double arena_bytes;
    int output_channels;
} opb200_net_info_t;

opb200_status opb200_net_create(opb200_net** out, const char* prototxt, const char* caffemodel, int gpu_id,
                                const char* last_blob_name);
void opb200_net_destroy(opb200_net* net);
opb200_status opb200_net_forward(opb200_net* net, const float* input, int input_on_device, const int shape[4],
                                 void* stream);
opb200_status opb200_net_output(opb200_net* net, const float** dev_ptr, int shape[4]);
/* The net's own fp32 NCHW input buffer for `shape` (re-plans buffers when the shape changes, like a forward): a producer
 * on the device (opb200_preprocess, opb200_crop_warp) writes the input there on `stream` and passes the same pointer to
 * opb200_net_forward(..., input_on_device = 1, ...), which then skips its input copy. No reference counterpart (Caffe's
 * input blob is written through mutable_gpu_data, netCaffe.cpp:230-236). */
opb200_status opb200_net_input_buffer(opb200_net* net, const int shape[4], float** dev_ptr, void* stream);
opb200_status opb200_net_info(opb200_net* net, opb200_net_info_t* info);
/* Test hook: any intermediate blob as fp32 NCHW on the host (host_nchw NULL = shape query). Synchronises. */
opb200_status opb200_net_read_blob(opb200_net* net, const char* name, float* host_nchw, size_t capacity,
                                   int shape[4]);

/* Host-only (no CUDA call): parse + plan the model and write one text line per planned kernel launch. */
opb200_status opb200_net_describe(const char* prototxt, const char* caffemodel, const char* last_blob_name,
                                  char* text, size_t capacity);
/* Host-only: the activation levels of the planned graph for an (n, 3, h, w) input -- size, number of convs reading the
 * level, their widest padding, and whether the level is stored frame-interleaved and with how many zero separator columns
 * (text lines "level L h= w= convs= max_pad= flat= sep="). No reference counterpart: it exposes the layout decision of
 * the reshape (reference: NetCaffe::reshapeNetCaffe, src/openpose/net/netCaffe.cpp:105-122) to tests without a GPU. */
opb200_status opb200_net_describe_shape(const char* prototxt, const char* caffemodel, const char* last_blob_name, int n,
                                        int h, int w, char* text, size_t capacity);

/* ---------------------------------------------------------------------------------------------------
 * Output serialisation -- host only, no CUDA call.
 * opb200_people_json / _save -- replace op::savePeopleJson and op::PeopleJsonSaver::save
 *   (reference: include/openpose/filestream/fileStream.hpp:37-45, src/openpose/filestream/fileStream.cpp:20-130
 *   addKeypointsToJson / addCandidatesToJson, :289-344 savePeopleJson, src/openpose/filestream/jsonOfstream.cpp,
 *   src/openpose/filestream/peopleJsonSaver.cpp:15-30; wire format doc/02_output.md, version 1.3). Each
 *   opb200_json_array is one (Array<float>, name) pair of the reference's keypointVector: `people` = getSize(0),
 *   `per_person` = getSize(1)*getSize(2) (0 for an empty Array, 1 for a 1-D array such as person_id).
 *   candidates: n_parts part lists, candidate_counts[part] triples (x, y, score) each, concatenated
 *   (PoseExtractorNet::getCandidatesCopy layout); n_parts = 0 omits "part_candidates" as an empty vector does.
 *   The bytes equal what the reference's std::ofstream receives (numbers as `ostream << float`, tabs and newlines
 *   only when human_readable). out = NULL with capacity 0 is a size query; *length is always set.
 * opb200_people_json_batch -- every frame of a connector batch (keypoints (frames,capacity,parts,3), counts) in one
 *   pass, each frame the document WPeopleJsonSaver writes for a body-only Datum
 *   (include/openpose/filestream/wPeopleJsonSaver.hpp:68-97): frame f = out[offsets[f], offsets[f+1]).
 * opb200_coco_json_* -- replace op::CocoJsonSaver (reference: include/openpose/filestream/cocoJsonSaver.hpp:16-40,
 *   src/openpose/filestream/cocoJsonSaver.cpp:28-277): open = constructor (same checks, same file names for the
 *   variants bit mask), record = CocoJsonSaver::record, close = destructor. format = op::CocoJsonFormat
 *   (0 body, 4 foot, 5 car); pose_model = op::PoseModel.
 * ------------------------------------------------------------------------------------------------- */
typedef struct opb200_json_array
{
    const char* name;
    const float* data;   /* host, row-major (people, per_person) */
    int people;
    int per_person;
} opb200_json_array;
opb200_status opb200_people_json(const opb200_json_array* arrays, int n_arrays, const float* candidates,
                                 const int* candidate_counts, int n_parts, int human_readable, char* out,
                                 size_t capacity, size_t* length);
opb200_status opb200_people_json_save(const char* path, const opb200_json_array* arrays, int n_arrays,
                                      const float* candidates, const int* candidate_counts, int n_parts,
                                      int human_readable);
opb200_status opb200_people_json_batch(const float* keypoints, const int* counts, int frames, int capacity, int parts,
                                       int human_readable, char* out, size_t out_capacity, size_t* offsets);
typedef struct opb200_coco_json opb200_coco_json;
opb200_status opb200_coco_json_open(const char* path, int pose_model, int human_readable, int variants, int format,
                                    int variant, opb200_coco_json** handle);
opb200_status opb200_coco_json_record(opb200_coco_json* handle, const float* keypoints, const float* scores,
                                      int people, int parts, const char* image_name,
                                      unsigned long long frame_number);
opb200_status opb200_coco_json_close(opb200_coco_json* handle);

/* ---------------------------------------------------------------------------------------------------
 * Multi-GPU body-pose engine: u8 BGR frames in, people out, frame order kept -- the hot path of the reference's
 * worker topology (one PoseExtractor per GPU on its own thread fed round-robin, results re-ordered by frame id:
 * include/openpose/wrapper/wrapperAuxiliary.hpp:181-219,326-338, include/openpose/thread/wQueueOrderer.hpp:12-80) and
 * of the per-frame chain CvMatToOpInput (device branch) -> PoseExtractorCaffe::forwardPass.
 * One host thread + compute stream + copy stream per device, `lanes` batches in flight per device, batches dealt to
 * the devices round-robin, no cross-device exchange. Per batch: H2D of the u8 frames, opb200_preprocess into the net's
 * input buffer, opb200_net_forward, opb200_resize_merge, opb200_nms, opb200_connect, D2H of the results.
 *   acquire -> pinned host staging buffer of batch * frame_h * frame_w * 3 bytes to write frames into (blocks while every
 *              lane of the next device is busy); commit(first_frame_id, n) -> enqueue its first n frames;
 *   submit  = acquire + memcpy + commit;
 *   next    -> the oldest submitted batch not yet returned (blocks until it is finished): keypoints (n, capacity, parts,
 *              3), scores (n, capacity), counts (n) into HOST buffers of at least `batch` frames.
 * acquire / commit / submit from one producer thread, next from one consumer thread (they may be the same).
 * ------------------------------------------------------------------------------------------------- */
typedef struct opb200_engine opb200_engine;
typedef struct opb200_engine_config
{
    int pose_model;                 /* OPB200_POSE_* */
    int net_w, net_h;               /* net input resolution, multiples of 8 (656 x 368) */
    int frame_w, frame_h;           /* producer resolution of the u8 BGR frames */
    int batch;                      /* frames per batch (per device) */
    int lanes;                      /* batches in flight per device, 1..8 (2 overlaps H2D with compute) */
    int capacity;                   /* people kept per frame */
    float scale_input_to_net;       /* ScaleAndSizeExtractor::extract's scaleInputToNetInputs[0] */
    float nms_threshold;            /* PoseProperty::NMSThreshold */
    opb200_connect_params connect;  /* incl. scale_factor = scaleNetToOutput and semantics (also used by nms) */
} opb200_engine_config;
opb200_status opb200_engine_create(opb200_engine** out, const char* prototxt, const char* caffemodel,
                                   const int* gpu_ids, int n_gpus, const opb200_engine_config* cfg);
void opb200_engine_destroy(opb200_engine* engine);
opb200_status opb200_engine_acquire(opb200_engine* engine, unsigned char** staging);
opb200_status opb200_engine_commit(opb200_engine* engine, unsigned long long first_frame_id, int n_frames);
opb200_status opb200_engine_submit(opb200_engine* engine, unsigned long long first_frame_id,
                                   const unsigned char* frames, int n_frames);
opb200_status opb200_engine_next(opb200_engine* engine, unsigned long long* first_frame_id, int* n_frames,
                                 float* keypoints, float* scores, int* counts);
int opb200_engine_in_flight(opb200_engine* engine);   /* batches submitted and not yet returned by next */

/* ---------------------------------------------------------------------------------------------------
 * Multi-GPU hand / face crop engine: the same worker topology for the keypoint nets that run on per-person crops
 * (BASELINE config 4: "per-person crops sharded across GPUs"). A batch = ONE u8 BGR frame + the 2x3 double inverse maps
 * (crop pixel -> frame position, what opb200_crop_warp takes) of up to `batch` crops of it: both hands of every person,
 * or every face. Per batch: H2D, opb200_crop_warp into the net's input buffer, opb200_net_forward, opb200_resize_merge
 * (x8), opb200_maximum, D2H -- the device part of HandExtractorCaffe::forwardPass / FaceExtractorCaffe::forwardPass
 * (reference: src/openpose/hand/handExtractorCaffe.cpp:305-464, src/openpose/face/faceExtractorCaffe.cpp:175-300) for all
 * crops of the batch at once. Batches go round the devices, results come back in submission order:
 *   submit(frame_id, frame, matrices, n)  (or opb200_engine_acquire: staging = frame bytes, then the matrices at
 *                                          opb200_crop_engine_matrix_offset(); opb200_engine_commit(frame_id, n));
 *   next -> peaks (n, heat_channels - 1, 3) = (x, y, value) of every part's maximum in CROP coordinates; the caller maps
 *           them back through its affine exactly as the reference does after maximumGpu.
 * Created with opb200_crop_engine_create, destroyed with opb200_engine_destroy; opb200_engine_in_flight applies.
 * ------------------------------------------------------------------------------------------------- */
typedef struct opb200_crop_engine_config
{
    int heat_channels;              /* channels of the net's output incl. background: 22 (hand), 71 (face) */
    int net_side;                   /* crop / net input side, a multiple of 8 (368) */
    int frame_w, frame_h;           /* resolution of the u8 BGR frames */
    int batch;                      /* crops per batch (per device) */
    int lanes;                      /* batches in flight per device, 1..8 */
} opb200_crop_engine_config;
opb200_status opb200_crop_engine_create(opb200_engine** out, const char* prototxt, const char* caffemodel,
                                        const int* gpu_ids, int n_gpus, const opb200_crop_engine_config* cfg);
size_t opb200_crop_engine_matrix_offset(opb200_engine* engine);   /* byte offset of the matrices in a staging buffer */
opb200_status opb200_crop_engine_submit(opb200_engine* engine, unsigned long long frame_id, const unsigned char* frame,
                                        const double* matrices, int n_crops);
opb200_status opb200_crop_engine_next(opb200_engine* engine, unsigned long long* frame_id, int* n_crops, float* peaks);

#ifdef __cplusplus
}
#endif
#endif /* OPB200_H */
