// loadProbe / loadColor (HelloPathtracing_sv4_vmv23/main.cpp:154-180) for the headless host.
// The reference reads its environment map with stbi_loadf(file, &w, &h, &channels, 4) from the vendored stb_image.h, which is not
// copied: readHDR below is our own reader of the published Radiance picture format (RGBE), keeping stbi_loadf's conventions:
//   * header "#?RADIANCE" or "#?RGBE", must contain FORMAT=32-bit_rle_rgbe, resolution line "-Y <h> +X <w>" only;
//   * top scanline first, no vertical flip; 4 channels out, alpha = 1;
//   * value = mantissa * 2^(exponent - 136); exponent 0 -> (0, 0, 0);
//   * widths 8..32767: new-style run-length scanlines (2, 2, w>>8, w&255, then 4 channel planes); if the FIRST scanline does not
//     start like that the whole file is read as flat RGBE quads, as it is for other widths.
// tests/golden/hdr.npz (minted with the reference's own stb_image) pins this reader bit for bit.
#include "../include/ref_api/Probe.h"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

inline void rgbeToFloat(const unsigned char* q, Color& out) {
    if (q[3] != 0) {
        const float f = std::ldexp(1.0f, (int)q[3] - 136);
        out = Color{q[0] * f, q[1] * f, q[2] * f, 1.0f};
    } else out = Color{0.f, 0.f, 0.f, 1.0f};
}

std::runtime_error bad(const std::string& file, const char* why) { return std::runtime_error("Could not load HDR probe " + file + ": " + why); }

}  // namespace

Color* readHDR(const std::string& file, int& width, int& height) {
    FILE* f = fopen(file.c_str(), "rb");
    if (!f) throw bad(file, "file not found");
    std::vector<unsigned char> buf;
    unsigned char chunk[1 << 16];
    for (size_t n; (n = fread(chunk, 1, sizeof chunk, f)) > 0;) buf.insert(buf.end(), chunk, chunk + n);
    fclose(f);
    size_t pos = 0;
    auto line = [&]() {
        std::string s;
        while (pos < buf.size() && buf[pos] != '\n') s.push_back((char)buf[pos++]);
        if (pos >= buf.size()) throw bad(file, "truncated header");
        ++pos;
        return s;
    };
    const std::string magic = line();
    if (magic != "#?RADIANCE" && magic != "#?RGBE") throw bad(file, "not a Radiance picture");
    bool valid = false;
    for (;;) {
        const std::string s = line();
        if (s.empty()) break;
        if (s == "FORMAT=32-bit_rle_rgbe") valid = true;
    }
    if (!valid) throw bad(file, "unsupported format");
    const std::string res = line();
    int w = 0, h = 0;
    if (sscanf(res.c_str(), "-Y %d +X %d", &h, &w) != 2 || w <= 0 || h <= 0) throw bad(file, "unsupported data layout");
    const unsigned char* p = buf.data() + pos;
    const unsigned char* end = buf.data() + buf.size();
    Color* out = new Color[(size_t)w * h];
    try {
        bool flat = w < 8 || w >= 32768;
        if (!flat && end - p >= 4 && (p[0] != 2 || p[1] != 2 || (p[2] & 0x80))) flat = true;
        if (flat) {
            if ((size_t)(end - p) < (size_t)4 * w * h) throw bad(file, "truncated data");
            for (size_t i = 0; i < (size_t)w * h; ++i) rgbeToFloat(p + 4 * i, out[i]);
        } else {
            std::vector<unsigned char> scan((size_t)4 * w);
            for (int j = 0; j < h; ++j) {
                if (end - p < 4 || p[0] != 2 || p[1] != 2 || (p[2] & 0x80) || ((p[2] << 8) | p[3]) != w) throw bad(file, "invalid decoded scanline length");
                p += 4;
                for (int k = 0; k < 4; ++k) {
                    int i = 0;
                    while (i < w) {
                        if (p >= end) throw bad(file, "truncated data");
                        int count = *p++;
                        if (count > 128) {
                            count -= 128;
                            if (count > w - i || p >= end) throw bad(file, "corrupt RLE data");
                            memset(&scan[(size_t)k * w + i], *p++, count);
                        } else {
                            if (count == 0 || count > w - i || end - p < count) throw bad(file, "corrupt RLE data");
                            memcpy(&scan[(size_t)k * w + i], p, count);
                            p += count;
                        }
                        i += count;
                    }
                }
                for (int i = 0; i < w; ++i) {
                    const unsigned char q[4] = {scan[i], scan[(size_t)w + i], scan[(size_t)2 * w + i], scan[(size_t)3 * w + i]};
                    rgbeToFloat(q, out[(size_t)j * w + i]);
                }
            }
        }
    } catch (...) { delete[] out;  throw; }
    width = w;  height = h;
    return out;
}

void loadProbe(ProbeData& probe, const std::string& hdrFile) {                                   // main.cpp:154-165
    probe.data = readHDR(hdrFile, probe.width, probe.height);
    probe.BuildCDF();
}

void loadColor(ProbeData& probe, const float3& color, int width, int height) {                   // main.cpp:167-180 (3840x2160 there)
    probe.width = width;  probe.height = height;
    probe.data = new Color[(size_t)width * height];
    for (size_t i = 0; i < (size_t)width * height; ++i) probe.data[i] = Color{color.x, color.y, color.z, 1.0f};
    probe.BuildCDF();
}
