/* A plain C99 consumer of libopb200.so: the frame-level engine from u8 frames in a file to people on stdout. Shows what a
 * host other than C++ / Python binds (include/opb200.h is strict C99) and is run by tests/test_c_example_gpu.py, which
 * compares its output with the Python mirror on the same frames.
 * usage: abi_example <prototxt> <caffemodel> <frames.raw> <n_frames> <frame_w> <frame_h> <net_w> <net_h> <scale_input_to_net>
 *                    <scale_net_to_output> <batch>                                                                         */
#include "opb200.h"

#include <stdio.h>
#include <stdlib.h>

static int fail(const char* what)
{
    fprintf(stderr, "%s: %s\n", what, opb200_last_error());
    return 1;
}

int main(int argc, char** argv)
{
    if (argc != 12)
    {
        fprintf(stderr, "usage: %s prototxt caffemodel frames.raw n w h net_w net_h scale_in scale_out batch\n", argv[0]);
        return 2;
    }
    const int n = atoi(argv[4]), w = atoi(argv[5]), h = atoi(argv[6]), batch = atoi(argv[11]);
    const int capacity = 64, parts = 25;
    const size_t frameBytes = (size_t)w * h * 3;
    opb200_engine_config cfg;
    opb200_engine* eng = NULL;
    int gpu = 0, done = 0, submitted = 0, rc = 0;
    unsigned char* frames = (unsigned char*)malloc(frameBytes * n);
    float* kp = (float*)malloc(sizeof(float) * batch * capacity * parts * 3);
    float* sc = (float*)malloc(sizeof(float) * batch * capacity);
    int* cnt = (int*)malloc(sizeof(int) * batch);
    FILE* f = fopen(argv[3], "rb");
    if (!frames || !kp || !sc || !cnt || !f || fread(frames, frameBytes, n, f) != (size_t)n)
    {
        fprintf(stderr, "cannot read %d frames from %s\n", n, argv[3]);
        return 2;
    }
    fclose(f);
    if (opb200_device_count() < 1)
        return fail("no CUDA device");
    cfg.pose_model = OPB200_POSE_BODY_25;
    cfg.net_w = atoi(argv[7]);
    cfg.net_h = atoi(argv[8]);
    cfg.frame_w = w;
    cfg.frame_h = h;
    cfg.batch = batch;
    cfg.lanes = 2;
    cfg.capacity = capacity;
    cfg.scale_input_to_net = (float)atof(argv[9]);
    cfg.nms_threshold = 0.02f;                     /* loose thresholds: a random-weight net emits noise */
    cfg.connect.pose_model = OPB200_POSE_BODY_25;
    cfg.connect.max_peaks = 127;
    cfg.connect.inter_min_above_threshold = 0.3f;
    cfg.connect.inter_threshold = 0.0f;
    cfg.connect.min_subset_cnt = 2;
    cfg.connect.min_subset_score = 0.0f;
    cfg.connect.default_nms_threshold = 0.02f;
    cfg.connect.scale_factor = (float)atof(argv[10]);
    cfg.connect.maximize_positives = 0;
    cfg.connect.semantics = OPB200_SEMANTICS_CPU;
    if (opb200_engine_create(&eng, argv[1], argv[2], &gpu, 1, &cfg) != OPB200_OK)
        return fail("opb200_engine_create");
    while (done < n && rc == 0)
    {
        /* keep up to `lanes` batches in flight, take the oldest back when both lanes are busy or everything is submitted */
        if (submitted < n && opb200_engine_in_flight(eng) < cfg.lanes)
        {
            const int k = n - submitted < batch ? n - submitted : batch;
            if (opb200_engine_submit(eng, (unsigned long long)submitted, frames + frameBytes * submitted, k) != OPB200_OK)
                rc = fail("opb200_engine_submit");
            submitted += k;
        }
        else
        {
            unsigned long long first = 0;
            int k = 0, i;
            if (opb200_engine_next(eng, &first, &k, kp, sc, cnt) != OPB200_OK)
            {
                rc = fail("opb200_engine_next");
                break;
            }
            for (i = 0; i < k; i++)
            {
                const int people = cnt[i] < capacity ? cnt[i] : capacity;
                const float* p0 = kp + (size_t)i * capacity * parts * 3;
                printf("frame %llu people %d", first + (unsigned long long)i, cnt[i]);
                if (people > 0)   /* first person: score and the hex bits of its first keypoint triple */
                    printf(" score %.9g kp %a %a %a", sc[(size_t)i * capacity], p0[0], p0[1], p0[2]);
                printf("\n");
            }
            done += k;
        }
    }
    opb200_engine_destroy(eng);
    free(frames);
    free(kp);
    free(sc);
    free(cnt);
    return rc;
}
