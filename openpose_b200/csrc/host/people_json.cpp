// Output serialisation of the sharded stream: the reference's per-frame people JSON
// (src/openpose/filestream/fileStream.cpp:20-130,289-344 addKeypointsToJson / addCandidatesToJson /
// savePeopleJson, driven by PeopleJsonSaver::save, peopleJsonSaver.cpp:15-30, and WPeopleJsonSaver,
// include/openpose/filestream/wPeopleJsonSaver.hpp:55-100) and the COCO results file
// (src/openpose/filestream/cocoJsonSaver.cpp:28-277), byte for byte, written into memory instead of through a
// std::ofstream per frame (jsonOfstream.cpp): one pass over the keypoints of a whole batch, one growing buffer,
// no stream objects, no locale machinery. Numbers are formatted exactly as `std::ostream << float` does
// (printf "%g": 6 significant digits) by an integer fast path that is exact for the fixed-notation range and
// falls back to snprintf elsewhere. Pure host code: no CUDA in this file.
#include "../../../include/opb200.h"
#include <cerrno>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace opb
{
    void setError(const char* fmt, ...);   // common.cu: thread-local text behind opb200_last_error()
}

namespace
{
    struct Out
    {
        std::string s;
        bool human;
        long long depth = 0;                       // braces + brackets, as JsonOfstream counts them (jsonOfstream.cpp:5-21)

        explicit Out(bool humanReadable) : human(humanReadable) {}
        void enterAndTab()
        {
            if (human)
            {
                s.push_back('\n');
                s.append((size_t)depth, '\t');
            }
        }
        void objectOpen() { ++depth; s.push_back('{'); }
        void objectClose() { --depth; enterAndTab(); s.push_back('}'); }
        void arrayOpen() { ++depth; s.push_back('['); enterAndTab(); }
        void arrayClose() { --depth; enterAndTab(); s.push_back(']'); }
        void key(const char* name) { enterAndTab(); s.push_back('"'); s.append(name); s.append("\":"); }
        void comma() { s.push_back(','); }
        void enter() { enterAndTab(); }
        void text(const char* t) { s.append(t); }
        void number(float v);
        void integer(unsigned long long v)
        {
            char b[24];
            const int n = std::snprintf(b, sizeof(b), "%llu", v);
            s.append(b, (size_t)n);
        }
    };

    // `ostream << float` == printf("%g"): P = 6 significant digits, fixed notation when the decimal exponent X of the
    // ROUNDED value satisfies -4 <= X < 6, trailing zeros (and a bare point) removed. Fast path for 1e-4 <= |v| < 1e6
    // v * 10^(5 - X) is formed EXACTLY in double (24-bit significand times 5^k, k <= 9, fits 53 bits), so the
    // round-half-even to six digits sees the same exact value glibc's arbitrary-precision printf does. Everything else
    // (exponent notation, NaN, inf) -> snprintf.
    void Out::number(float v)
    {
        char b[40];
        if (v == 0.f)
        {
            s.append(std::signbit(v) ? "-0" : "0");
            return;
        }
        const float a = std::fabs(v);
        if (!(a >= 1e-4f && a < 999999.4f))        // also NaN / inf; 999999.5 would round up to 1e+06 (exponent form)
        {
            const int n = std::snprintf(b, sizeof(b), "%g", (double)v);
            s.append(b, (size_t)n);
            return;
        }
        static const double p10[] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9};
        const double d = (double)a;
        int X = d >= 1e5 ? 5 : d >= 1e4 ? 4 : d >= 1e3 ? 3 : d >= 1e2 ? 2 : d >= 1e1 ? 1 : d >= 1e0 ? 0
                : d >= 1e-1 ? -1 : d >= 1e-2 ? -2 : d >= 1e-3 ? -3 : -4;
        // six significant digits as an integer in [100000, 999999], round half to even like glibc
        double scaled = d * p10[5 - X];            // exact: see above
        long long digits = std::llrint(scaled);    // FE_TONEAREST: ties to even
        if (digits >= 1000000)                     // 9.999995 -> 10.0000: one more integer digit
        {
            ++X;
            digits = 100000;
        }
        char* p = b;
        if (std::signbit(v))
            *p++ = '-';
        char dg[6];
        for (int i = 5; i >= 0; --i)
        {
            dg[i] = (char)('0' + digits % 10);
            digits /= 10;
        }
        int last = 5;                              // strip trailing zeros of the fraction
        while (last > 0 && dg[last] == '0')
            --last;
        if (X >= 0)
        {
            for (int i = 0; i <= X; ++i)
                *p++ = dg[i];
            if (last > X)
            {
                *p++ = '.';
                for (int i = X + 1; i <= last; ++i)
                    *p++ = dg[i];
            }
        }
        else
        {
            *p++ = '0';
            *p++ = '.';
            for (int i = -1; i > X; --i)
                *p++ = '0';
            for (int i = 0; i <= last; ++i)
                *p++ = dg[i];
        }
        s.append(b, (size_t)(p - b));
    }

    // addKeypointsToJson, fileStream.cpp:20-78
    void addKeypoints(Out& o, const opb200_json_array* arrays, int nArrays)
    {
        o.key("people");
        o.arrayOpen();
        int numberPeople = 0;
        for (int v = 0; v < nArrays; ++v)
            numberPeople = arrays[v].people > numberPeople ? arrays[v].people : numberPeople;
        for (int person = 0; person < numberPeople; ++person)
        {
            o.objectOpen();
            for (int v = 0; v < nArrays; ++v)
            {
                const opb200_json_array& a = arrays[v];
                const int perRow = a.per_person;
                o.key(a.name);
                o.arrayOpen();
                if (perRow > 0)
                {
                    const float* row = a.data + (size_t)person * perRow;
                    for (int e = 0; e < perRow - 1; ++e)
                    {
                        o.number(row[e]);
                        o.comma();
                    }
                    o.number(row[perRow - 1]);
                }
                o.arrayClose();
                if (v < nArrays - 1)
                    o.comma();
            }
            o.objectClose();
            if (person < numberPeople - 1)
            {
                o.comma();
                o.enter();
            }
        }
        o.arrayClose();
    }

    // addCandidatesToJson, fileStream.cpp:80-128
    void addCandidates(Out& o, const float* cand, const int* counts, int nParts)
    {
        o.key("part_candidates");
        o.arrayOpen();
        o.objectOpen();
        const float* p = cand;
        for (int part = 0; part < nParts; ++part)
        {
            char name[16];
            std::snprintf(name, sizeof(name), "%d", part);
            o.key(name);
            o.arrayOpen();
            const int n = counts[part];
            for (int c = 0; c < n; ++c, p += 3)
            {
                o.number(p[0]);
                o.comma();
                o.number(p[1]);
                o.comma();
                o.number(p[2]);
                if (c < n - 1)
                    o.comma();
            }
            o.arrayClose();
            if (part < nParts - 1)
                o.comma();
        }
        o.objectClose();
        o.arrayClose();
    }

    // savePeopleJson, fileStream.cpp:306-344 (+ the final enterAndTab of ~JsonOfstream, jsonOfstream.cpp:75-83)
    void peopleJson(Out& o, const opb200_json_array* arrays, int nArrays, const float* cand, const int* counts,
                    int nParts)
    {
        o.objectOpen();
        o.key("version");
        o.text("1.3");
        o.comma();
        addKeypoints(o, arrays, nArrays);
        if (nParts > 0)
        {
            o.comma();
            addCandidates(o, cand, counts, nParts);
        }
        o.objectClose();
        o.enterAndTab();
    }

    opb200_status fail(opb200_status st, const char* msg)
    {
        opb::setError("%s", msg);
        return st;
    }

    opb200_status checkArrays(const opb200_json_array* arrays, int nArrays)
    {
        if (nArrays < 0 || (nArrays > 0 && arrays == nullptr))
            return fail(OPB200_ERR_INVALID, "people json: null array list");
        for (int v = 0; v < nArrays; ++v)
        {
            if (arrays[v].name == nullptr || arrays[v].people < 0 || arrays[v].per_person < 0)
                return fail(OPB200_ERR_INVALID, "people json: array without a name or with negative sizes");
            if (arrays[v].people > 0 && arrays[v].per_person > 0 && arrays[v].data == nullptr)
                return fail(OPB200_ERR_INVALID, "people json: null data for a non-empty array");
        }
        return OPB200_OK;
    }

    opb200_status deliver(const std::string& s, char* out, size_t capacity, size_t* length)
    {
        if (length != nullptr)
            *length = s.size();
        if (out == nullptr || capacity < s.size())
            return out == nullptr && capacity == 0 ? OPB200_OK   // size query
                                                   : fail(OPB200_ERR_INVALID, "people json: output buffer too small");
        std::memcpy(out, s.data(), s.size());
        return OPB200_OK;
    }

    opb200_status writeFile(const char* path, const std::string& s, const char* mode)
    {
        std::FILE* f = std::fopen(path, mode);
        if (f == nullptr)   // jsonOfstream.cpp:31-32
            return fail(OPB200_ERR_IO, (std::string("Json file ") + path + " could not be opened.").c_str());
        const size_t n = std::fwrite(s.data(), 1, s.size(), f);
        const int rc = std::fclose(f);
        if (n != s.size() || rc != 0)
            return fail(OPB200_ERR_IO, (std::string("Json file ") + path + " could not be written.").c_str());
        return OPB200_OK;
    }
}

extern "C" opb200_status opb200_people_json(const opb200_json_array* arrays, int n_arrays, const float* candidates,
                                            const int* candidate_counts, int n_parts, int human_readable, char* out,
                                            size_t capacity, size_t* length)
{
    const opb200_status st = checkArrays(arrays, n_arrays);
    if (st != OPB200_OK)
        return st;
    if (n_parts < 0 || (n_parts > 0 && (candidate_counts == nullptr)))
        return fail(OPB200_ERR_INVALID, "people json: candidates without counts");
    Out o(human_readable != 0);
    peopleJson(o, arrays, n_arrays, candidates, candidate_counts, n_parts);
    return deliver(o.s, out, capacity, length);
}

extern "C" opb200_status opb200_people_json_save(const char* path, const opb200_json_array* arrays, int n_arrays,
                                                 const float* candidates, const int* candidate_counts, int n_parts,
                                                 int human_readable)
{
    if (path == nullptr)
        return fail(OPB200_ERR_INVALID, "people json: null path");
    const opb200_status st = checkArrays(arrays, n_arrays);
    if (st != OPB200_OK)
        return st;
    if (n_parts < 0 || (n_parts > 0 && (candidate_counts == nullptr)))
        return fail(OPB200_ERR_INVALID, "people json: candidates without counts");
    Out o(human_readable != 0);
    peopleJson(o, arrays, n_arrays, candidates, candidate_counts, n_parts);
    return writeFile(path, o.s, "wb");
}

// A whole batch in the connector's output layout: frame f holds counts[f] people at keypoints + f*capacity*parts*3.
// Emits, per frame, what WPeopleJsonSaver writes for a body-only run (person_id = -1 per person, empty face / hand /
// 3-D arrays; wPeopleJsonSaver.hpp:75-93); frame f occupies out[offsets[f], offsets[f+1]).
extern "C" opb200_status opb200_people_json_batch(const float* keypoints, const int* counts, int frames, int capacity,
                                                  int parts, int human_readable, char* out, size_t out_capacity,
                                                  size_t* offsets)
{
    if (frames < 0 || capacity < 0 || parts < 0 || (frames > 0 && (keypoints == nullptr || counts == nullptr)) ||
        offsets == nullptr)
        return fail(OPB200_ERR_INVALID, "people json batch: invalid argument");
    static const char* kNames[9] = {"person_id", "pose_keypoints_2d", "face_keypoints_2d", "hand_left_keypoints_2d",
                                    "hand_right_keypoints_2d", "pose_keypoints_3d", "face_keypoints_3d",
                                    "hand_left_keypoints_3d", "hand_right_keypoints_3d"};
    Out o(human_readable != 0);
    std::vector<float> ids;
    offsets[0] = 0;
    for (int f = 0; f < frames; ++f)
    {
        const int people = counts[f];
        if (people < 0 || people > capacity)
            return fail(OPB200_ERR_INVALID, "people json batch: people count outside [0, capacity]");
        ids.assign((size_t)people, -1.f);
        opb200_json_array a[9];
        for (int v = 0; v < 9; ++v)
            a[v] = opb200_json_array{kNames[v], nullptr, 0, 0};
        a[0] = opb200_json_array{kNames[0], ids.data(), people, people > 0 ? 1 : 0};
        a[1] = opb200_json_array{kNames[1], keypoints + (size_t)f * capacity * parts * 3, people,
                                 people > 0 ? parts * 3 : 0};
        o.depth = 0;
        peopleJson(o, a, 9, nullptr, nullptr, 0);
        offsets[f + 1] = o.s.size();
    }
    if (out == nullptr && out_capacity == 0)
        return OPB200_OK;   // size query: offsets[frames] is the total
    if (out == nullptr || out_capacity < o.s.size())
        return fail(OPB200_ERR_INVALID, "people json batch: output buffer too small");
    std::memcpy(out, o.s.data(), o.s.size());
    return OPB200_OK;
}

// ----------------------------------------------------------------------------------------------------
// COCO results file (cocoJsonSaver.cpp). One object per valid person, streamed into the files the
// reference opens: <path> (body or car), <path>_foot.<ext>, ... (only the body and foot variants exist
// for the models on this path: BODY_25 / COCO_18; the others need the 135-part model).
// ----------------------------------------------------------------------------------------------------
struct opb200_coco_json
{
    struct Stream
    {
        std::string path;
        int format;               // 0 body, 1 hand21, 2 hand42, 3 face, 4 foot, 5 car (CocoJsonFormat, filestream/enumClasses.hpp)
        bool firstAdded = false;
        Out out;
        Stream(const std::string& p, int f, bool human) : path(p), format(f), out(human) {}
    };
    std::vector<Stream> streams;
    int poseModel;
    int variant;
};

namespace
{
    // getLastNumber, utilities/string.cpp:8-20
    bool lastNumber(const std::string& s, unsigned long long& v)
    {
        const std::string digits = s.substr(s.find_last_not_of("0123456789") + 1);
        if (digits.empty())
            return false;
        errno = 0;
        char* end = nullptr;
        v = std::strtoull(digits.c_str(), &end, 10);
        return errno == 0 && end != digits.c_str();
    }
}

extern "C" opb200_status opb200_coco_json_open(const char* path, int pose_model, int human_readable, int variants,
                                               int format, int variant, opb200_coco_json** handle)
{
    if (handle == nullptr)
        return fail(OPB200_ERR_INVALID, "coco json: null handle");
    *handle = nullptr;
    if (path == nullptr || path[0] == '\0')   // cocoJsonSaver.cpp:37-39
        return fail(OPB200_ERR_INVALID, "Empty path given as output file path for saving COCO JSON format.");
    if (variants >= 32)                       // :40-42
        return fail(OPB200_ERR_INVALID, "Unknown value for cocoJsonFormat (flag `--write_coco_json_variants`).");
    const std::string full(path);
    const size_t dot = full.find_last_of('.');                       // getFullFilePathNoExtension
    const std::string noExt = dot != std::string::npos ? full.substr(0, dot) : full;
    const size_t slash = full.find_last_of("/\\");                    // getFileExtension works on name + extension
    const std::string nameExt = slash != std::string::npos ? full.substr(slash + 1) : full;
    const size_t dot2 = nameExt.find_last_of('.');
    const std::string ext = dot2 != std::string::npos ? nameExt.substr(dot2 + 1) : "";
    auto* h = new opb200_coco_json();
    h->poseModel = pose_model;
    h->variant = variant;
    const bool human = human_readable != 0;
    if (variants % 2 == 1 || variants < 1)
        h->streams.emplace_back(full, format, human);
    if ((variants / 2) % 2 == 1 || variants < 1)
        h->streams.emplace_back(noExt + "_foot." + ext, 4, human);
    if ((variants / 4) % 2 == 1 || variants < 1)
        h->streams.emplace_back(noExt + "_face." + ext, 3, human);
    if ((variants / 8) % 2 == 1 || variants < 1)
        h->streams.emplace_back(noExt + "_hand21." + ext, 1, human);
    if ((variants / 16) % 2 == 1 || variants < 1)
        h->streams.emplace_back(noExt + "_hand42." + ext, 2, human);
    for (auto& s : h->streams)
    {
        const opb200_status st = writeFile(s.path.c_str(), std::string(), "wb");   // JsonOfstream opens (truncates) now
        if (st != OPB200_OK)
        {
            delete h;
            return st;
        }
        s.out.arrayOpen();
    }
    *handle = h;
    return OPB200_OK;
}

extern "C" opb200_status opb200_coco_json_record(opb200_coco_json* h, const float* keypoints, const float* scores,
                                                 int people, int parts, const char* image_name,
                                                 unsigned long long frame_number)
{
    if (h == nullptr || people < 0 || parts < 0 || (people > 0 && (keypoints == nullptr || scores == nullptr)))
        return fail(OPB200_ERR_INVALID, "coco json: invalid argument");
    if (people == 0)
        return OPB200_OK;
    const std::string name(image_name != nullptr ? image_name : "");
    for (auto& s : h->streams)
    {
        std::vector<int> idx;
        unsigned long long imageId = frame_number;
        const bool needsId = s.format == 0 || s.format == 4 || s.format == 5;
        if (needsId && !lastNumber(name, imageId))
            return fail(OPB200_ERR_INVALID,
                        (std::string("`--write_coco_json` is to be used with the original ") +
                         (s.format == 5 ? "car" : "COCO") +
                         " dataset images. If you are not applying those, OpenPose cannot obtain the ID from their file"
                         " names. Image name:\n" + name).c_str());
        if (needsId)
            imageId = (unsigned long long)(int)imageId;   // the reference narrows to int (cocoJsonSaver.cpp:13)
        if (s.format == 0)
        {
            if (parts == 23)
                idx = {0, 14, 13, 16, 15, 4, 1, 5, 2, 6, 3, 10, 7, 11, 8, 12, 9};
            else if (parts == 18)
                idx = {0, 15, 14, 17, 16, 5, 2, 6, 3, 7, 4, 11, 8, 12, 9, 13, 10};
            else if (h->poseModel == 13 || h->poseModel == 14)   // BODY_25B, BODY_135 (pose/enumClasses.hpp:27-28)
            {
                idx.resize(17);
                for (int i = 0; i < 17; ++i)
                    idx[i] = i;
            }
            else if (parts == 19 || parts == 25 || parts == 59)
                idx = {0, 16, 15, 18, 17, 5, 2, 6, 3, 7, 4, 12, 9, 13, 10, 14, 11};
        }
        else if (s.format == 4)
        {
            if (parts == 25 || parts > 60)
                idx = {19, 20, 21, 22, 23, 24};
            else if (parts == 23)
                idx = {17, 18, 19, 20, 21, 22};
        }
        else if (s.format == 5)
        {
            if (parts == 12)
                idx = {0, 1, 2, 3, 4, 5, 6, 7, 8, 8, 9, 10, 11, 11};
            else if (parts == 22)
            {
                if (h->variant == 0)
                    idx = {0, 1, 2, 3, 6, 7, 12, 13, 14, 15, 16, 17};
                else if (h->variant == 1)
                    idx = {0, 1, 2, 3, 6, 7, 12, 13, 14, 15, 20, 21};
                else if (h->variant == 2)
                    for (int i = 0; i < 20; ++i)
                        idx.push_back(i);
            }
        }
        // face / hand21 / hand42 need the 135-part model (cocoJsonSaver.cpp:147-177): out of scope -> same error
        if (idx.empty())
            return fail(OPB200_ERR_INVALID,
                        (std::string("Invalid number of body parts (") + std::to_string(parts) + ").").c_str());
        Out& o = s.out;
        for (int person = 0; person < people; ++person)
        {
            bool any = false;
            for (size_t b = 0; b < idx.size() && !any; ++b)
                any = keypoints[3 * ((size_t)person * parts + idx[b]) + 2] > 0.f;
            if (!any)
                continue;
            if (s.firstAdded)
            {
                o.comma();
                o.enter();
            }
            else
                s.firstAdded = true;
            o.objectOpen();
            o.key("image_id");
            o.integer(imageId);
            o.comma();
            o.key("category_id");
            o.text("1");
            o.comma();
            o.key("keypoints");
            o.arrayOpen();
            for (size_t b = 0; b < idx.size(); ++b)
            {
                const float* kp = keypoints + 3 * ((size_t)person * parts + idx[b]);
                const bool valid = kp[2] > 0.f;
                o.number(valid ? kp[0] : -1.f);
                o.comma();
                o.number(valid ? kp[1] : -1.f);
                o.comma();
                o.text(valid ? "1" : "0");
                if (b + 1 < idx.size())
                    o.comma();
            }
            o.arrayClose();
            o.comma();
            o.key("score");
            o.number(scores[person]);
            o.objectClose();
        }
        if (o.s.size() > (1u << 20))   // flush: keep the buffer small on long streams
        {
            const opb200_status st = writeFile(s.path.c_str(), o.s, "ab");
            if (st != OPB200_OK)
                return st;
            o.s.clear();
        }
    }
    return OPB200_OK;
}

extern "C" opb200_status opb200_coco_json_close(opb200_coco_json* h)
{
    if (h == nullptr)
        return OPB200_OK;
    opb200_status st = OPB200_OK;
    for (auto& s : h->streams)
    {
        s.out.arrayClose();       // ~CocoJsonSaver, cocoJsonSaver.cpp:80-90
        s.out.enterAndTab();      // ~JsonOfstream
        const opb200_status w = writeFile(s.path.c_str(), s.out.s, "ab");
        if (w != OPB200_OK)
            st = w;
    }
    delete h;
    return st;
}
