// See caffe_format.h. Written against the protobuf text/wire format specifications; semantics of the fields
// follow the reference's caffe.proto (paths in the header).
#include "caffe_format.h"
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

namespace opb
{
    bool readFile(const std::string& path, std::string& out)
    {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f)
            return false;
        std::fseek(f, 0, SEEK_END);
        const long n = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        out.resize(n > 0 ? (size_t)n : 0);
        const size_t got = n > 0 ? std::fread(&out[0], 1, (size_t)n, f) : 0;
        std::fclose(f);
        return got == out.size();
    }

    // ------------------------------------------------------------------------------------------------
    // text format: generic tree first (field name -> scalar string | sub-message), then typed extraction
    // ------------------------------------------------------------------------------------------------
    namespace
    {
        struct Node;
        struct Field
        {
            std::string name;
            std::string scalar;                 // valid when !msg
            std::shared_ptr<Node> msg;
        };
        struct Node
        {
            std::vector<Field> fields;
            std::vector<const Field*> all(const char* n) const
            {
                std::vector<const Field*> r;
                for (const auto& f : fields)
                    if (f.name == n)
                        r.push_back(&f);
                return r;
            }
            const Field* first(const char* n) const
            {
                for (const auto& f : fields)
                    if (f.name == n)
                        return &f;
                return nullptr;
            }
            std::string str(const char* n, const std::string& d = "") const
            {
                const Field* f = first(n);
                return (f && !f->msg) ? f->scalar : d;
            }
            int integer(const char* n, int d) const
            {
                const Field* f = first(n);
                return (f && !f->msg) ? std::atoi(f->scalar.c_str()) : d;
            }
        };

        struct Lexer
        {
            const std::string& s;
            size_t i = 0;
            explicit Lexer(const std::string& text) : s(text) {}
            void skip()
            {
                while (i < s.size())
                {
                    const char c = s[i];
                    if (c == '#')
                        while (i < s.size() && s[i] != '\n')
                            ++i;
                    else if (c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == ',' || c == ';')
                        ++i;
                    else
                        break;
                }
            }
            bool eof()
            {
                skip();
                return i >= s.size();
            }
            char peek()
            {
                skip();
                return i < s.size() ? s[i] : '\0';
            }
            bool token(std::string& out, std::string& err)
            {
                skip();
                out.clear();
                if (i >= s.size())
                {
                    err = "prototxt: unexpected end of file";
                    return false;
                }
                const char c = s[i];
                if (c == '"' || c == '\'')
                {
                    ++i;
                    while (i < s.size() && s[i] != c)
                    {
                        if (s[i] == '\\' && i + 1 < s.size())
                            ++i;
                        out.push_back(s[i++]);
                    }
                    if (i >= s.size())
                    {
                        err = "prototxt: unterminated string";
                        return false;
                    }
                    ++i;
                    return true;
                }
                if (c == '{' || c == '}' || c == '<' || c == '>' || c == ':')
                {
                    out.push_back(c);
                    ++i;
                    return true;
                }
                while (i < s.size())
                {
                    const char d = s[i];
                    if (d == ' ' || d == '\t' || d == '\n' || d == '\r' || d == ':' || d == '{' || d == '}' ||
                        d == '<' || d == '>' || d == '#' || d == ',' || d == ';' || d == '"')
                        break;
                    out.push_back(d);
                    ++i;
                }
                if (out.empty())
                {
                    err = "prototxt: cannot tokenize";
                    return false;
                }
                return true;
            }
        };

        constexpr int kMaxNesting = 64;   // a prototxt nests 3-4 deep; a file of repeated '{' must not overflow the stack

        bool parseNode(Lexer& lx, Node& node, char closer, std::string& err, int depth = 0)
        {
            if (depth > kMaxNesting)
            {
                err = "prototxt: messages nested deeper than 64 levels";
                return false;
            }
            for (;;)
            {
                if (lx.eof())
                {
                    if (closer)
                        err = "prototxt: missing closing brace";
                    return closer == '\0';
                }
                if (closer && lx.peek() == closer)
                {
                    std::string t;
                    return lx.token(t, err);
                }
                Field f;
                if (!lx.token(f.name, err))
                    return false;
                if (lx.peek() == ':')
                {
                    std::string t;
                    lx.token(t, err);
                }
                const char c = lx.peek();
                if (c == '{' || c == '<')
                {
                    std::string t;
                    lx.token(t, err);
                    f.msg = std::make_shared<Node>();
                    if (!parseNode(lx, *f.msg, c == '{' ? '}' : '>', err, depth + 1))
                        return false;
                }
                else if (!lx.token(f.scalar, err))
                    return false;
                node.fields.push_back(std::move(f));
            }
        }

        bool truthy(const std::string& s, bool d)
        {
            if (s.empty())
                return d;
            return s == "true" || s == "True" || s == "1" || s == "t";
        }
    }

    bool parsePrototxt(const std::string& text, NetDef& out, std::string& err)
    {
        Lexer lx(text);
        Node root;
        if (!parseNode(lx, root, '\0', err))
            return false;
        out = NetDef();
        out.name = root.str("name");
        for (const Field* f : root.all("input"))
            out.inputs.push_back(f->scalar);
        for (const Field* f : root.all("input_dim"))
            out.inputDim.push_back(std::atoi(f->scalar.c_str()));
        for (const Field* sh : root.all("input_shape"))   // legacy alternative to input_dim
            if (sh->msg)
                for (const Field* d : sh->msg->all("dim"))
                    out.inputDim.push_back(std::atoi(d->scalar.c_str()));
        if (!root.all("layers").empty())
        {
            err = "prototxt: V1 'layers' definitions are not supported (upgrade the file to 'layer')";
            return false;
        }
        for (const Field* lf : root.all("layer"))
        {
            if (!lf->msg)
                continue;
            const Node& L = *lf->msg;
            LayerDef d;
            d.name = L.str("name");
            d.type = L.str("type");
            for (const Field* f : L.all("bottom"))
                d.bottoms.push_back(f->scalar);
            for (const Field* f : L.all("top"))
                d.tops.push_back(f->scalar);
            if (const Field* cp = L.first("convolution_param"))
                if (cp->msg)
                {
                    const Node& c = *cp->msg;
                    d.numOutput = c.integer("num_output", 0);
                    d.kernel = c.integer("kernel_size", c.integer("kernel_h", 1));
                    d.pad = c.integer("pad", c.integer("pad_h", 0));
                    d.stride = c.integer("stride", c.integer("stride_h", 1));
                    d.group = c.integer("group", 1);
                    d.dilation = c.integer("dilation", 1);
                    d.biasTerm = truthy(c.str("bias_term"), true);
                    if (c.integer("kernel_w", d.kernel) != d.kernel || c.integer("pad_w", d.pad) != d.pad ||
                        c.integer("stride_w", d.stride) != d.stride)
                    {
                        err = "prototxt: non-square convolution in layer " + d.name;
                        return false;
                    }
                }
            if (const Field* pp = L.first("pooling_param"))
                if (pp->msg)
                {
                    const Node& c = *pp->msg;
                    const std::string m = c.str("pool", "MAX");
                    d.poolMethod = (m == "MAX" || m == "0") ? 0 : ((m == "AVE" || m == "1") ? 1 : 2);
                    d.kernel = c.integer("kernel_size", c.integer("kernel_h", 1));
                    d.stride = c.integer("stride", c.integer("stride_h", 1));
                    d.pad = c.integer("pad", c.integer("pad_h", 0));
                    if (c.integer("kernel_w", d.kernel) != d.kernel || c.integer("pad_w", d.pad) != d.pad ||
                        c.integer("stride_w", d.stride) != d.stride)
                    {
                        err = "prototxt: non-square pooling in layer " + d.name;
                        return false;
                    }
                }
            if (const Field* cc = L.first("concat_param"))
                if (cc->msg)
                    d.axis = cc->msg->integer("axis", cc->msg->integer("concat_dim", 1));
            if (const Field* pr = L.first("prelu_param"))
                if (pr->msg)
                    d.channelShared = truthy(pr->msg->str("channel_shared"), false);
            if (const Field* ip = L.first("input_param"))
                if (ip->msg)
                    if (const Field* sh = ip->msg->first("shape"))
                        if (sh->msg)
                            for (const Field* dm : sh->msg->all("dim"))
                                d.inputShape.push_back(std::atoi(dm->scalar.c_str()));
            out.layers.push_back(std::move(d));
        }
        return true;
    }

    // ------------------------------------------------------------------------------------------------
    // wire format
    // ------------------------------------------------------------------------------------------------
    namespace
    {
        struct Reader
        {
            const uint8_t* p;
            const uint8_t* end;
            bool ok = true;
            uint64_t varint()
            {
                uint64_t v = 0;
                int shift = 0;
                while (p < end && shift < 64)
                {
                    const uint8_t b = *p++;
                    v |= (uint64_t)(b & 0x7F) << shift;
                    if (!(b & 0x80))
                        return v;
                    shift += 7;
                }
                ok = false;
                return 0;
            }
            // Reads one field header; for length-delimited fields sets sub to the payload.
            bool next(uint32_t& field, uint32_t& wire, Reader& sub, uint64_t& value)
            {
                if (p >= end || !ok)
                    return false;
                const uint64_t key = varint();
                if (!ok)
                    return false;
                field = (uint32_t)(key >> 3);
                wire = (uint32_t)(key & 7);
                switch (wire)
                {
                case 0: value = varint(); return ok;
                case 1:
                    if (end - p < 8) { ok = false; return false; }
                    std::memcpy(&value, p, 8);
                    p += 8;
                    return true;
                case 2:
                {
                    const uint64_t n = varint();
                    if (!ok || (uint64_t)(end - p) < n) { ok = false; return false; }
                    sub.p = p;
                    sub.end = p + n;
                    sub.ok = true;
                    p += n;
                    return true;
                }
                case 5:
                    if (end - p < 4) { ok = false; return false; }
                    value = 0;
                    std::memcpy(&value, p, 4);
                    p += 4;
                    return true;
                default: ok = false; return false;
                }
            }
        };

        bool parseBlob(Reader r, BlobData& b)
        {
            int legacy[4] = {-1, -1, -1, -1};
            uint32_t f, w;
            Reader sub{nullptr, nullptr};
            uint64_t v;
            std::vector<double> dbl;
            while (r.next(f, w, sub, v))
            {
                if (f >= 1 && f <= 4 && w == 0)
                    legacy[f - 1] = (int)v;
                else if (f == 5 && w == 2)   // packed float data
                {
                    const size_t n = (size_t)(sub.end - sub.p) / 4;
                    const size_t old = b.data.size();
                    b.data.resize(old + n);
                    std::memcpy(b.data.data() + old, sub.p, n * 4);
                }
                else if (f == 5 && w == 5)   // unpacked float
                {
                    float x;
                    const uint32_t u = (uint32_t)v;
                    std::memcpy(&x, &u, 4);
                    b.data.push_back(x);
                }
                else if (f == 8 && w == 2)   // packed double_data
                {
                    const size_t n = (size_t)(sub.end - sub.p) / 8;
                    for (size_t i = 0; i < n; ++i)
                    {
                        double x;
                        std::memcpy(&x, sub.p + 8 * i, 8);
                        dbl.push_back(x);
                    }
                }
                else if (f == 7 && w == 2)   // BlobShape { repeated int64 dim = 1 [packed] }
                {
                    Reader s = sub;
                    uint32_t f2, w2;
                    Reader sub2{nullptr, nullptr};
                    uint64_t v2;
                    while (s.next(f2, w2, sub2, v2))
                    {
                        if (f2 == 1 && w2 == 2)
                            while (sub2.p < sub2.end && sub2.ok)
                                b.shape.push_back((int)sub2.varint());
                        else if (f2 == 1 && w2 == 0)
                            b.shape.push_back((int)v2);
                    }
                    if (!s.ok)
                        return false;
                }
            }
            if (!r.ok)
                return false;
            if (b.data.empty() && !dbl.empty())
                b.data.assign(dbl.begin(), dbl.end());
            if (b.shape.empty() && legacy[0] >= 0)
                b.shape.assign(legacy, legacy + 4);
            return true;
        }
    }

    bool readCaffemodel(const std::string& path, WeightMap& out, std::string& err)
    {
        std::string buf;
        if (!readFile(path, buf))
        {
            err = "caffemodel: cannot read " + path;
            return false;
        }
        Reader r{reinterpret_cast<const uint8_t*>(buf.data()), reinterpret_cast<const uint8_t*>(buf.data()) + buf.size()};
        uint32_t f, w;
        Reader sub{nullptr, nullptr};
        uint64_t v;
        while (r.next(f, w, sub, v))
        {
            if (f == 2 && w == 2)
            {
                err = "caffemodel: V1 'layers' weights are not supported";
                return false;
            }
            if (f != 100 || w != 2)
                continue;
            Reader L = sub;
            std::string name;
            std::vector<BlobData> blobs;
            uint32_t f2, w2;
            Reader sub2{nullptr, nullptr};
            uint64_t v2;
            while (L.next(f2, w2, sub2, v2))
            {
                if (f2 == 1 && w2 == 2)
                    name.assign(reinterpret_cast<const char*>(sub2.p), (size_t)(sub2.end - sub2.p));
                else if (f2 == 7 && w2 == 2)
                {
                    BlobData b;
                    if (!parseBlob(sub2, b))
                    {
                        err = "caffemodel: malformed blob in layer " + name;
                        return false;
                    }
                    blobs.push_back(std::move(b));
                }
            }
            if (!L.ok)
            {
                err = "caffemodel: malformed layer record";
                return false;
            }
            if (!blobs.empty())
                out[name] = std::move(blobs);
        }
        if (!r.ok)
        {
            err = "caffemodel: malformed file " + path;
            return false;
        }
        return true;
    }
}
