// vrl_io.cpp -- see vrl_io.h.
#include "vrl_io.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <stdexcept>

namespace vrlhost {
namespace {
template <typename T> T rd(std::ifstream &f, const std::string &path) {
    T v;
    f.read(reinterpret_cast<char *>(&v), sizeof(T));
    if (!f) throw std::runtime_error("Unexpected end of volume file " + path);
    return v;
}
}  // namespace

VolumeGrid read_vol(const std::string &path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("Volume file \"" + path + "\" not found!");
    char header[3] = { 0, 0, 0 };
    f.read(header, 3);
    if (!f || header[0] != 'V' || header[1] != 'O' || header[2] != 'L') throw std::runtime_error("Invalid volume file " + path);
    int version = rd<uint8_t>(f, path);
    if (version != 3) throw std::runtime_error("Invalid version, currently only version 3 is supported (found " + std::to_string(version) + ")");
    int32_t type = rd<int32_t>(f, path);
    if (type != 1) throw std::runtime_error("Wrong type, currently only type == 1 (Float32) data is supported (found type = " + std::to_string(type) + ")");
    VolumeGrid g;
    g.nx = rd<int32_t>(f, path); g.ny = rd<int32_t>(f, path); g.nz = rd<int32_t>(f, path);
    if (g.nx < 1 || g.ny < 1 || g.nz < 1 || (int64_t) g.nx * g.ny * g.nz < 8)
        throw std::runtime_error("Invalid grid dimensions: " + std::to_string(g.nx) + " x " + std::to_string(g.ny) + " x " + std::to_string(g.nz) +
                                 " < 8 (must have at least one value at each corner)");
    g.channels = rd<int32_t>(f, path);
    if (g.channels < 1 || g.channels > 4) throw std::runtime_error("Invalid channel count in volume file " + path);
    for (int k = 0; k < 3; ++k) g.bbox_min[k] = rd<float>(f, path);
    for (int k = 0; k < 3; ++k) g.bbox_max[k] = rd<float>(f, path);
    size_t n = (size_t) g.nx * g.ny * g.nz * (size_t) g.channels;
    g.data.resize(n);
    f.read(reinterpret_cast<char *>(g.data.data()), (std::streamsize) (n * sizeof(float)));
    if (!f) throw std::runtime_error("Unexpected end of volume file " + path);
    g.max = -std::numeric_limits<float>::infinity(); g.mean = 0.0;
    for (float v : g.data) { g.mean += (double) v; g.max = std::max(g.max, v); }
    g.mean /= (double) n;
    return g;
}

void write_vol(const std::string &path, const VolumeGrid &g) {
    if ((size_t) g.nx * g.ny * g.nz * (size_t) g.channels != g.data.size()) throw std::runtime_error("write_vol: data size does not match the dimensions");
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot write " + path);
    const char hdr[4] = { 'V', 'O', 'L', 3 };
    f.write(hdr, 4);
    int32_t type = 1;
    f.write(reinterpret_cast<const char *>(&type), 4);
    f.write(reinterpret_cast<const char *>(&g.nx), 4); f.write(reinterpret_cast<const char *>(&g.ny), 4); f.write(reinterpret_cast<const char *>(&g.nz), 4);
    f.write(reinterpret_cast<const char *>(&g.channels), 4);
    f.write(reinterpret_cast<const char *>(g.bbox_min), 12); f.write(reinterpret_cast<const char *>(g.bbox_max), 12);
    f.write(reinterpret_cast<const char *>(g.data.data()), (std::streamsize) (g.data.size() * sizeof(float)));
}

std::vector<float> read_obj(const std::string &path, const float *to_world) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("OBJ file \"" + path + "\" not found!");
    std::vector<float> pos, tris;
    std::string line;
    size_t lineno = 0;
    auto emit = [&](size_t i) {
        float p[3] = { pos[i * 3], pos[i * 3 + 1], pos[i * 3 + 2] };
        if (to_world) {
            float q[4];
            for (int r = 0; r < 4; ++r) q[r] = to_world[r * 4] * p[0] + to_world[r * 4 + 1] * p[1] + to_world[r * 4 + 2] * p[2] + to_world[r * 4 + 3];
            float iw = q[3] != 0.f ? 1.f / q[3] : 1.f;
            p[0] = q[0] * iw; p[1] = q[1] * iw; p[2] = q[2] * iw;
        }
        tris.insert(tris.end(), p, p + 3);
    };
    while (std::getline(f, line)) {
        ++lineno;
        const char *cur = line.c_str();
        while (*cur == ' ' || *cur == '\t') ++cur;
        if (cur[0] == 'v' && (cur[1] == ' ' || cur[1] == '\t')) {
            char *end;
            float p[3];
            cur += 2;
            for (int k = 0; k < 3; ++k) { p[k] = std::strtof(cur, &end); if (end == cur) throw std::runtime_error(path + ":" + std::to_string(lineno) + ": could not parse vertex"); cur = end; }
            if (!std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2])) throw std::runtime_error(path + ": mesh contains invalid vertex position data");
            pos.insert(pos.end(), p, p + 3);
        } else if (cur[0] == 'f' && (cur[1] == ' ' || cur[1] == '\t')) {
            cur += 2;
            size_t first = 0, prev = 0, count = 0;
            while (true) {
                char *end;
                unsigned long v = std::strtoul(cur, &end, 10);      // the reference parses with strtoul too (obj.cpp:214)
                if (end == cur) break;
                cur = end;
                while (*cur == '/') { ++cur; (void) std::strtoul(cur, &end, 10); cur = end; }   // skip vt / vn indices
                size_t idx = (size_t) v - 1;
                if (v == 0 || idx >= pos.size() / 3) throw std::runtime_error(path + ": reference to invalid vertex " + std::to_string(v) + "!");
                if (count == 0) first = idx;
                else if (count >= 2) { emit(first); emit(prev); emit(idx); }      // fan, as obj.cpp:256-262
                prev = idx; ++count;
                while (*cur == ' ' || *cur == '\t' || *cur == '\r') ++cur;
            }
            if (count < 3) throw std::runtime_error(path + ":" + std::to_string(lineno) + ": face with fewer than 3 vertices");
        }
    }
    return tris;
}

std::vector<float> develop_rgb(const std::vector<float> &xyzaw, int width, int height) {
    if (xyzaw.size() != (size_t) width * height * 5) throw std::runtime_error("develop_rgb: XYZAW buffer size mismatch");
    std::vector<float> rgb((size_t) width * height * 3, 0.f);
    for (size_t i = 0; i < (size_t) width * height; ++i) {
        const float *p = &xyzaw[i * 5];
        if (p[4] == 0.f) continue;
        float iw = 1.f / p[4], X = p[0] * iw, Y = p[1] * iw, Z = p[2] * iw;
        rgb[i * 3 + 0] = 3.240479f * X - 1.537150f * Y - 0.498535f * Z;      // xyz_to_srgb, spectrum.h:229-235
        rgb[i * 3 + 1] = -0.969256f * X + 1.875991f * Y + 0.041556f * Z;
        rgb[i * 3 + 2] = 0.055648f * X - 0.204043f * Y + 1.057311f * Z;
    }
    return rgb;
}

void write_pfm(const std::string &path, const std::vector<float> &rgb, int width, int height) {
    if (rgb.size() != (size_t) width * height * 3) throw std::runtime_error("write_pfm: RGB buffer size mismatch");
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot write " + path);
    std::ostringstream hdr;
    hdr << "PF\n" << width << " " << height << "\n-1.0\n";                 // negative scale = little endian
    f << hdr.str();
    for (int y = height - 1; y >= 0; --y)                                  // PFM stores the bottom row first
        f.write(reinterpret_cast<const char *>(&rgb[(size_t) y * width * 3]), (std::streamsize) ((size_t) width * 3 * sizeof(float)));
}

namespace {
void put_i32(std::string &b, int32_t v) { b.append(reinterpret_cast<const char *>(&v), 4); }
void put_f32(std::string &b, float v) { b.append(reinterpret_cast<const char *>(&v), 4); }
void put_str(std::string &b, const char *s) { b.append(s); b.push_back('\0'); }
// IEEE binary32 -> binary16, round to nearest even (what OpenEXR's `half` does)
uint16_t to_half(float f) {
    uint32_t x; std::memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u, mant = x & 0x7fffffu;
    const int32_t e = (int32_t) ((x >> 23) & 0xffu) - 127 + 15;
    if (((x >> 23) & 0xffu) == 0xffu) return (uint16_t) (sign | 0x7c00u | (mant ? 0x200u : 0u));     // inf / nan
    if (e >= 31) return (uint16_t) (sign | 0x7c00u);                                                    // overflow -> inf
    if (e <= 0) {                                                                                        // subnormal half or zero
        if (e < -10) return (uint16_t) sign;
        const uint32_t m = mant | 0x800000u; const int shift = 14 - e;
        uint32_t h = m >> shift; const uint32_t rem = m & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
        if (rem > halfway || (rem == halfway && (h & 1u))) ++h;
        return (uint16_t) (sign | h);
    }
    uint32_t h = ((uint32_t) e << 10) | (mant >> 13);
    const uint32_t rem = mant & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1u))) ++h;                                              // may carry into the exponent: still correct
    return (uint16_t) (sign | h);
}
void attr(std::string &b, const char *name, const char *type, const std::string &value) {
    put_str(b, name); put_str(b, type); put_i32(b, (int32_t) value.size()); b.append(value);
}
}  // namespace

// OpenEXR 2 single-part scan-line layout: magic, version, attributes (name, type, size, value), a zero byte, one 64-bit
// file offset per scan line, then per line { y, byte count, the line's channels in alphabetical order, each `width` floats }
void write_exr(const std::string &path, const std::vector<float> &xyzaw, int width, int height, bool half) {
    if (xyzaw.size() != (size_t) width * height * 5) throw std::runtime_error("write_exr: XYZAW buffer size mismatch");
    const std::vector<float> rgb = develop_rgb(xyzaw, width, height);
    std::string h;
    put_i32(h, 20000630); put_i32(h, 2);
    { std::string v; const char *names[4] = { "A", "B", "G", "R" };
      for (const char *n : names) { put_str(v, n); put_i32(v, half ? 1 /* HALF */ : 2 /* FLOAT */); v.append(4, '\0'); put_i32(v, 1); put_i32(v, 1); }
      v.push_back('\0'); attr(h, "channels", "chlist", v); }
    attr(h, "compression", "compression", std::string(1, '\0'));
    { std::string v; put_i32(v, 0); put_i32(v, 0); put_i32(v, width - 1); put_i32(v, height - 1);
      attr(h, "dataWindow", "box2i", v); attr(h, "displayWindow", "box2i", v); }
    attr(h, "lineOrder", "lineOrder", std::string(1, '\0'));
    { std::string v; put_f32(v, 1.f); attr(h, "pixelAspectRatio", "float", v); }
    { std::string v; put_f32(v, 0.f); put_f32(v, 0.f); attr(h, "screenWindowCenter", "v2f", v); }
    { std::string v; put_f32(v, 1.f); attr(h, "screenWindowWidth", "float", v); }
    h.push_back('\0');
    const uint64_t line_bytes = (uint64_t) width * 4 * (half ? 2 : 4), block = 8 + line_bytes;
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot write " + path);
    f.write(h.data(), (std::streamsize) h.size());
    for (int y = 0; y < height; ++y) { const uint64_t off = h.size() + (uint64_t) height * 8 + (uint64_t) y * block; f.write(reinterpret_cast<const char *>(&off), 8); }
    std::vector<float> line((size_t) width * 4);
    for (int y = 0; y < height; ++y) {
        for (int x = 0; x < width; ++x) {
            const size_t i = (size_t) y * width + x;
            const float w = xyzaw[i * 5 + 4];
            line[(size_t) x] = w != 0.f ? xyzaw[i * 5 + 3] / w : 0.f;                         // A
            line[(size_t) width + x] = rgb[i * 3 + 2];                                           // B
            line[(size_t) 2 * width + x] = rgb[i * 3 + 1];                                       // G
            line[(size_t) 3 * width + x] = rgb[i * 3 + 0];                                       // R
        }
        const int32_t yy = y, n = (int32_t) line_bytes;
        f.write(reinterpret_cast<const char *>(&yy), 4); f.write(reinterpret_cast<const char *>(&n), 4);
        if (half) {
            std::vector<uint16_t> hl(line.size());
            for (size_t k = 0; k < line.size(); ++k) hl[k] = to_half(line[k]);
            f.write(reinterpret_cast<const char *>(hl.data()), (std::streamsize) line_bytes);
        } else f.write(reinterpret_cast<const char *>(line.data()), (std::streamsize) line_bytes);
    }
}

void write_film(const std::string &out, const std::vector<float> &xyzaw, int width, int height, bool half) {
    const std::string ext = out.size() > 4 ? out.substr(out.size() - 4) : std::string();
    if (ext == ".exr") write_exr(out, xyzaw, width, height, half);
    else if (ext == ".pfm") write_pfm(out, develop_rgb(xyzaw, width, height), width, height);
    else { std::ofstream o(out, std::ios::binary); if (!o) throw std::runtime_error("cannot write " + out);
           o.write(reinterpret_cast<const char *>(xyzaw.data()), (std::streamsize) (xyzaw.size() * sizeof(float))); }
}

}  // namespace vrlhost
