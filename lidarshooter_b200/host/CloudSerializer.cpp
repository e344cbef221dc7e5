#include "CloudSerializer.hpp"

#include <cstdio>
#include <cstring>

namespace lidarshooter {

namespace {

void put32(std::vector<uint8_t> &o, uint32_t v)
{
    for (int k = 0; k < 4; ++k) o.push_back(static_cast<uint8_t>(v >> (8 * k)));
}
void putString(std::vector<uint8_t> &o, const std::string &s)
{
    put32(o, static_cast<uint32_t>(s.size()));
    o.insert(o.end(), s.begin(), s.end());
}

struct Reader {
    const uint8_t *p;
    size_t left;
    bool ok = true;
    uint32_t u32()
    {
        if (left < 4) { ok = false; return 0; }
        const uint32_t v = uint32_t(p[0]) | uint32_t(p[1]) << 8 | uint32_t(p[2]) << 16 | uint32_t(p[3]) << 24;
        p += 4; left -= 4;
        return v;
    }
    uint8_t u8()
    {
        if (left < 1) { ok = false; return 0; }
        const uint8_t v = *p;
        ++p; --left;
        return v;
    }
    bool bytes(size_t n, const uint8_t *&out)
    {
        if (left < n) { ok = false; return false; }
        out = p; p += n; left -= n;
        return true;
    }
    std::string str()
    {
        const uint32_t n = u32();
        const uint8_t *b = nullptr;
        if (!ok || !bytes(n, b)) return std::string();
        return std::string(reinterpret_cast<const char *>(b), n);
    }
};

}  // namespace

size_t serializedLength(const PointCloud2 &c)
{
    size_t n = 4 + 8 + 4 + c.header.frame_id.size() + 8 + 4;
    for (const auto &f : c.fields) n += 4 + f.name.size() + 4 + 1 + 4;
    return n + 1 + 4 + 4 + 4 + c.data.size() + 1;
}

size_t serializePointCloud2(const PointCloud2 &c, std::vector<uint8_t> &out, bool fixRowStep)
{
    const size_t start = out.size();
    out.reserve(start + serializedLength(c));
    put32(out, c.header.seq);
    put32(out, c.header.stamp.sec);
    put32(out, c.header.stamp.nsec);
    putString(out, c.header.frame_id);
    put32(out, c.height);
    put32(out, c.width);
    put32(out, static_cast<uint32_t>(c.fields.size()));
    for (const auto &f : c.fields) {
        putString(out, f.name);
        put32(out, f.offset);
        out.push_back(f.datatype);
        put32(out, f.count);
    }
    out.push_back(c.is_bigendian ? 1 : 0);
    put32(out, c.point_step);
    put32(out, fixRowStep ? c.width * c.point_step : c.row_step);
    put32(out, static_cast<uint32_t>(c.data.size()));
    out.insert(out.end(), c.data.begin(), c.data.end());
    out.push_back(c.is_dense ? 1 : 0);
    return out.size() - start;
}

size_t deserializePointCloud2(const uint8_t *buf, size_t len, PointCloud2 &c)
{
    Reader r{buf, len};
    c.header.seq = r.u32();
    const uint32_t sec = r.u32(), nsec = r.u32();
    c.header.stamp = Time(sec, nsec);
    c.header.frame_id = r.str();
    c.height = r.u32();
    c.width = r.u32();
    const uint32_t nf = r.u32();
    if (!r.ok || nf > r.left / 13) return 0;          // every field takes at least 13 bytes
    c.fields.assign(nf, PointField());
    for (auto &f : c.fields) {
        f.name = r.str();
        f.offset = r.u32();
        f.datatype = r.u8();
        f.count = r.u32();
    }
    c.is_bigendian = r.u8() != 0;
    c.point_step = r.u32();
    c.row_step = r.u32();
    const uint32_t nd = r.u32();
    const uint8_t *d = nullptr;
    if (!r.ok || !r.bytes(nd, d)) return 0;
    c.data.assign(d, d + nd);
    c.is_dense = r.u8() != 0;
    return r.ok ? len - r.left : 0;
}

bool appendCloudToFile(const std::string &path, const PointCloud2 &cloud, bool fixRowStep)
{
    std::vector<uint8_t> msg;
    put32(msg, 0);
    const size_t n = serializePointCloud2(cloud, msg, fixRowStep);
    for (int k = 0; k < 4; ++k) msg[k] = static_cast<uint8_t>(uint32_t(n) >> (8 * k));
    FILE *f = std::fopen(path.c_str(), "ab");
    if (!f) return false;
    const bool ok = std::fwrite(msg.data(), 1, msg.size(), f) == msg.size();
    return std::fclose(f) == 0 && ok;
}

}  // namespace lidarshooter
