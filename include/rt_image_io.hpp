// rt_image_io.hpp -- headless result sinks replacing the Direct2D preview (spheres-app.cpp:217-219; app.cpp:78-110):
//   PPM  (P6)  from m_backbufferLdr  (uint32 0xAARRGGBB as XMStoreColor packs it, app.h:27)
//   PFM  (PF)  linear mean radiance  (m_backbufferHdr sum / sample count, app.h:26)
//   EXR        OpenEXR 2 single-part scanline file, NO_COMPRESSION, FLOAT channels A,B,G,R holding the running SUM
//              (A = samples accumulated) + an int attribute "rtSampleCount": a render checkpoint that any EXR viewer opens.
// Header-only, no dependencies (OpenEXR / tinyexr are not in this image).  The reader only accepts files of the layout
// the writer produces (uncompressed, FLOAT, increasing y).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace rtio {

inline void write_all(const std::string& path, const std::vector<unsigned char>& bytes) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot open " + path + " for writing");
    const size_t n = std::fwrite(bytes.data(), 1, bytes.size(), f);
    const int rc = std::fclose(f);
    if (n != bytes.size() || rc != 0) throw std::runtime_error("short write to " + path);
}
inline std::vector<unsigned char> read_all(const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    std::vector<unsigned char> out;
    unsigned char buf[1 << 16];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) out.insert(out.end(), buf, buf + n);
    std::fclose(f);
    return out;
}

// bgra8: W*H uint32 0xAARRGGBB, row 0 = top
inline void write_ppm(const std::string& path, const uint32_t* bgra8, int w, int h) {
    char hdr[64];
    const int hl = std::snprintf(hdr, sizeof hdr, "P6\n%d %d\n255\n", w, h);
    std::vector<unsigned char> out(hl + static_cast<size_t>(w) * h * 3);
    std::memcpy(out.data(), hdr, hl);
    unsigned char* p = out.data() + hl;
    for (size_t i = 0; i < static_cast<size_t>(w) * h; ++i) {
        const uint32_t c = bgra8[i];
        *p++ = static_cast<unsigned char>((c >> 16) & 0xFF); *p++ = static_cast<unsigned char>((c >> 8) & 0xFF); *p++ = static_cast<unsigned char>(c & 0xFF);
    }
    write_all(path, out);
}

// hdr_sum: W*H float4 running sum; writes sum/sample_count as little-endian colour PFM (rows bottom-to-top per the format)
inline void write_pfm(const std::string& path, const float* hdr_sum, int w, int h, uint64_t sample_count) {
    char hdr[64];
    const int hl = std::snprintf(hdr, sizeof hdr, "PF\n%d %d\n-1.0\n", w, h);
    std::vector<unsigned char> out(hl + static_cast<size_t>(w) * h * 12);
    std::memcpy(out.data(), hdr, hl);
    float* p = reinterpret_cast<float*>(out.data() + hl);   // hl is not 4-aligned in general: go through memcpy
    std::vector<float> row(static_cast<size_t>(w) * 3);
    const float n = static_cast<float>(sample_count ? sample_count : 1);
    unsigned char* dst = out.data() + hl;
    (void)p;
    for (int y = h - 1; y >= 0; --y) {
        const float* src = hdr_sum + static_cast<size_t>(y) * w * 4;
        for (int x = 0; x < w; ++x) { row[3 * x] = src[4 * x] / n; row[3 * x + 1] = src[4 * x + 1] / n; row[3 * x + 2] = src[4 * x + 2] / n; }
        std::memcpy(dst, row.data(), row.size() * 4);
        dst += row.size() * 4;
    }
    write_all(path, out);
}

// ---- OpenEXR (scanline, uncompressed) ----
namespace detail {
struct Buf {
    std::vector<unsigned char> b;
    void raw(const void* p, size_t n) { const unsigned char* c = static_cast<const unsigned char*>(p); b.insert(b.end(), c, c + n); }
    void str(const char* s) { raw(s, std::strlen(s) + 1); }
    void i32(int32_t v) { raw(&v, 4); }
    void u64(uint64_t v) { raw(&v, 8); }
    void f32(float v) { raw(&v, 4); }
    void u8(unsigned char v) { b.push_back(v); }
    void attr(const char* name, const char* type, int32_t size) { str(name); str(type); i32(size); }
};
}  // namespace detail

inline void write_exr_sum(const std::string& path, const float* hdr_sum, int w, int h, uint64_t sample_count) {
    detail::Buf o;
    const unsigned char magic[8] = {0x76, 0x2f, 0x31, 0x01, 2, 0, 0, 0};
    o.raw(magic, 8);
    const char* names[4] = {"A", "B", "G", "R"};                 // alphabetical, as the format requires
    o.attr("channels", "chlist", 4 * (2 + 4 + 4 + 4 + 4) + 1);
    for (const char* n : names) { o.str(n); o.i32(2 /*FLOAT*/); o.u8(1 /*pLinear*/); o.u8(0); o.u8(0); o.u8(0); o.i32(1); o.i32(1); }
    o.u8(0);
    o.attr("compression", "compression", 1); o.u8(0);            // NO_COMPRESSION
    o.attr("dataWindow", "box2i", 16); o.i32(0); o.i32(0); o.i32(w - 1); o.i32(h - 1);
    o.attr("displayWindow", "box2i", 16); o.i32(0); o.i32(0); o.i32(w - 1); o.i32(h - 1);
    o.attr("lineOrder", "lineOrder", 1); o.u8(0);                // INCREASING_Y
    o.attr("pixelAspectRatio", "float", 4); o.f32(1.0f);
    o.attr("rtSampleCount", "int", 4); o.i32(static_cast<int32_t>(sample_count));
    o.attr("screenWindowCenter", "v2f", 8); o.f32(0.f); o.f32(0.f);
    o.attr("screenWindowWidth", "float", 4); o.f32(1.0f);
    o.u8(0);                                                     // end of header
    const size_t table = o.b.size();
    const size_t line_bytes = static_cast<size_t>(w) * 16;
    const size_t block = 8 + line_bytes;
    for (int y = 0; y < h; ++y) o.u64(table + static_cast<size_t>(h) * 8 + static_cast<size_t>(y) * block);
    const int src_lane[4] = {3, 2, 1, 0};                        // A,B,G,R <- float4 (r,g,b,w)
    std::vector<float> plane(static_cast<size_t>(w));
    o.b.reserve(o.b.size() + static_cast<size_t>(h) * block);
    for (int y = 0; y < h; ++y) {
        o.i32(y); o.i32(static_cast<int32_t>(line_bytes));
        const float* src = hdr_sum + static_cast<size_t>(y) * w * 4;
        for (int c = 0; c < 4; ++c) {
            for (int x = 0; x < w; ++x) plane[x] = src[4 * x + src_lane[c]];
            o.raw(plane.data(), plane.size() * 4);
        }
    }
    write_all(path, o.b);
}

// Reads a file written by write_exr_sum back into a float4 running sum.  Throws on any other EXR layout.
inline void read_exr_sum(const std::string& path, std::vector<float>& hdr_sum, int& w, int& h, uint64_t& sample_count) {
    const std::vector<unsigned char> b = read_all(path);
    size_t p = 8;
    auto need = [&](size_t n) { if (p + n > b.size()) throw std::runtime_error(path + ": truncated EXR"); };
    if (b.size() < 8 || b[0] != 0x76 || b[1] != 0x2f || b[2] != 0x31 || b[3] != 0x01 || b[4] != 2) throw std::runtime_error(path + ": not an OpenEXR 2 file");
    auto cstr = [&]() { std::string s; for (;;) { need(1); const char c = static_cast<char>(b[p++]); if (!c) break; s.push_back(c); } return s; };
    auto i32 = [&]() { need(4); int32_t v; std::memcpy(&v, &b[p], 4); p += 4; return v; };
    w = h = -1; sample_count = 0;
    bool ok_comp = false, ok_ch = false, ok_order = false;
    for (;;) {
        const std::string name = cstr();
        if (name.empty()) break;
        const std::string type = cstr();
        const int32_t size = i32();
        if (size < 0) throw std::runtime_error(path + ": negative attribute size");
        need(static_cast<size_t>(size));
        const size_t v = p;
        auto at_least = [&](int32_t n) { if (size < n) throw std::runtime_error(path + ": attribute '" + name + "' is shorter than its type"); };
        if (name == "dataWindow") at_least(16);
        else if (name == "compression" || name == "lineOrder") at_least(1);
        else if (name == "rtSampleCount") at_least(4);
        if (name == "dataWindow") { int32_t q[4]; std::memcpy(q, &b[v], 16); if (q[0] != 0 || q[1] != 0) throw std::runtime_error(path + ": dataWindow origin must be 0,0"); w = q[2] + 1; h = q[3] + 1; }
        else if (name == "compression") ok_comp = b[v] == 0;
        else if (name == "lineOrder") ok_order = b[v] == 0;
        else if (name == "rtSampleCount") { int32_t n; std::memcpy(&n, &b[v], 4); sample_count = static_cast<uint64_t>(n); }
        else if (name == "channels") {
            static const unsigned char expect_names[4] = {'A', 'B', 'G', 'R'};
            ok_ch = size == 4 * 18 + 1;
            for (int c = 0; ok_ch && c < 4; ++c) { int32_t t; std::memcpy(&t, &b[v + c * 18 + 2], 4); ok_ch = b[v + c * 18] == expect_names[c] && b[v + c * 18 + 1] == 0 && t == 2; }
        }
        p = v + static_cast<size_t>(size);
    }
    if (!ok_comp || !ok_ch || !ok_order || w <= 0 || h <= 0) throw std::runtime_error(path + ": not an rt_b200 checkpoint EXR (need uncompressed FLOAT A,B,G,R, increasing y)");
    p += static_cast<size_t>(h) * 8;                             // offset table
    hdr_sum.assign(static_cast<size_t>(w) * h * 4, 0.f);
    const int dst_lane[4] = {3, 2, 1, 0};
    for (int y = 0; y < h; ++y) {
        const int32_t yy = i32(), nbytes = i32();
        if (yy != y || nbytes != w * 16) throw std::runtime_error(path + ": unexpected scanline block");
        need(static_cast<size_t>(nbytes));
        for (int c = 0; c < 4; ++c) {
            const unsigned char* src = &b[p + static_cast<size_t>(c) * w * 4];
            float* dst = hdr_sum.data() + static_cast<size_t>(y) * w * 4 + dst_lane[c];
            for (int x = 0; x < w; ++x) std::memcpy(dst + 4 * x, src + 4 * x, 4);
        }
        p += static_cast<size_t>(nbytes);
    }
}

}  // namespace rtio
