// imageio.cpp -- HDR image readers for the `envmap` emitter: Portable Float Map and Radiance RGBE (.hdr / .rgbe).
// Reference: Bitmap::readPFM (src/libcore/bitmap.cpp:3764-3814) and Bitmap::readRGBE (:3590-3690, RGBE_ToFloat :3522-3530).
// OpenEXR / PNG / JPEG need libraries that are not on this box; asking for them is reported as an error.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>

#include "plugins.h"

namespace mtsbhost {

namespace {

std::string pfmToken(std::istream &is) {          // bitmap.cpp: pfmReadString -- characters up to the next whitespace
    std::string s;
    for (;;) {
        int c = is.get();
        if (c == EOF) error("readPFM(): unexpected end of file!");
        if (c == ' ' || c == '\t' || c == '\n' || c == '\r') { if (s.empty()) continue; break; }
        s.push_back((char) c);
    }
    return s;
}

void readPFM(std::istream &is, Image &img) {
    char header[3];
    is.read(header, 3);
    if (header[0] != 'P' || !(header[1] == 'F' || header[1] == 'f')) error("readPFM(): Invalid header!");
    const bool color = header[1] == 'F';
    char *end = nullptr;
    const std::string ws = pfmToken(is);
    img.width = (int) strtol(ws.c_str(), &end, 10);
    if (*end != '\0') error("Could not parse image dimensions!");
    const std::string hs = pfmToken(is);
    img.height = (int) strtol(hs.c_str(), &end, 10);
    if (*end != '\0') error("Could not parse image dimensions!");
    const std::string ss = pfmToken(is);
    const float scaleAndOrder = (float) strtod(ss.c_str(), &end);
    if (*end != '\0') error("Could not parse scale/order information!");
    if (img.width <= 0 || img.height <= 0) error("readPFM(): invalid image size!");
    const size_t ch = color ? 3 : 1, count = (size_t) img.width * img.height * ch;
    std::vector<float> raw(count);
    is.read(reinterpret_cast<char *>(raw.data()), (std::streamsize) (count * sizeof(float)));
    if ((size_t) is.gcount() != count * sizeof(float)) error("readPFM(): unexpected end of file!");
    if (scaleAndOrder > 0.0f) {                    // big endian payload
        for (float &v : raw) { unsigned char *b = reinterpret_cast<unsigned char *>(&v); std::swap(b[0], b[3]); std::swap(b[1], b[2]); }
    }
    const float scale = std::fabs(scaleAndOrder);
    if (scale != 1) for (float &v : raw) v *= scale;
    // the file stores the bottom row first (flipVertically, bitmap.cpp:3813)
    img.rgb.resize((size_t) img.width * img.height * 3);
    for (int y = 0; y < img.height; ++y)
        for (int x = 0; x < img.width; ++x)
            for (int c = 0; c < 3; ++c)
                img.rgb[3 * ((size_t) y * img.width + x) + c] = raw[ch * ((size_t) (img.height - 1 - y) * img.width + x) + (color ? c : 0)];
}

inline void rgbeToFloat(const unsigned char rgbe[4], float *out) {      // bitmap.cpp:3522-3530
    if (rgbe[3]) {
        const float f = std::ldexp(1.0f, (int) rgbe[3] - (128 + 8));
        for (int i = 0; i < 3; ++i) out[i] = rgbe[i] * f;
    } else out[0] = out[1] = out[2] = 0.0f;
}

void readFlat(std::istream &is, float *data, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        unsigned char rgbe[4];
        is.read(reinterpret_cast<char *>(rgbe), 4);
        if (is.gcount() != 4) error("readRGBE(): unexpected end of file!");
        rgbeToFloat(rgbe, data + 3 * i);
    }
}

void readRGBE(std::istream &is, Image &img) {
    std::string line;
    std::getline(is, line);
    if (line.length() < 2 || line[0] != '#' || line[1] != '?') error("readRGBE(): Invalid header!");
    bool recognized = false;
    for (;;) {
        if (!std::getline(is, line)) error("readRGBE(): parser error!");
        if (line.compare(0, 22, "FORMAT=32-bit_rle_rgbe") == 0) recognized = true;
        if (line.compare(0, 3, "-Y ") == 0) {
            if (sscanf(line.c_str(), "-Y %i +X %i", &img.height, &img.width) < 2) error("readRGBE(): parser error!");
            break;
        }
    }
    if (!recognized) error("readRGBE(): invalid format!");
    if (img.width <= 0 || img.height <= 0) error("readRGBE(): invalid image size!");
    img.rgb.assign((size_t) img.width * img.height * 3, 0.0f);
    float *data = img.rgb.data();
    if (img.width < 8 || img.width > 0x7fff) { readFlat(is, data, (size_t) img.width * img.height); return; }
    std::vector<unsigned char> buffer((size_t) 4 * img.width);
    for (int y = 0; y < img.height; ++y) {
        unsigned char rgbe[4];
        is.read(reinterpret_cast<char *>(rgbe), 4);
        if (is.gcount() != 4) error("readRGBE(): unexpected end of file!");
        if (rgbe[0] != 2 || rgbe[1] != 2 || (rgbe[2] & 0x80)) {
            // not run-length encoded (only legal on the first scanline, as in the reference)
            rgbeToFloat(rgbe, data);
            readFlat(is, data + 3, (size_t) img.width * img.height - 1 - (size_t) y * img.width);
            return;
        }
        if ((((int) rgbe[2]) << 8 | rgbe[3]) != img.width) error("readRGBE(): wrong scanline width!");
        unsigned char *ptr = buffer.data();
        for (int i = 0; i < 4; ++i) {
            unsigned char *ptr_end = buffer.data() + (size_t) (i + 1) * img.width;
            while (ptr < ptr_end) {
                unsigned char buf[2];
                is.read(reinterpret_cast<char *>(buf), 2);
                if (is.gcount() != 2) error("readRGBE(): unexpected end of file!");
                if (buf[0] > 128) {
                    int count = buf[0] - 128;
                    if (count == 0 || count > ptr_end - ptr) error("readRGBE(): bad scanline data!");
                    while (count-- > 0) *ptr++ = buf[1];
                } else {
                    int count = buf[0];
                    if (count == 0 || count > ptr_end - ptr) error("readRGBE(): bad scanline data!");
                    *ptr++ = buf[1];
                    if (--count > 0) { is.read(reinterpret_cast<char *>(ptr), count); if (is.gcount() != count) error("readRGBE(): unexpected end of file!"); }
                    ptr += count;
                }
            }
        }
        for (int i = 0; i < img.width; ++i) {
            const unsigned char px[4] = {buffer[i], buffer[(size_t) img.width + i], buffer[(size_t) 2 * img.width + i], buffer[(size_t) 3 * img.width + i]};
            rgbeToFloat(px, data);
            data += 3;
        }
    }
}

}  // namespace

Image readImage(const std::string &path) {
    std::ifstream is(path, std::ios::binary);
    if (!is) error("Environment map file \"" + path + "\" could not be found!");          // envmap.cpp:121-122
    char magic[2] = {0, 0};
    is.read(magic, 2);
    is.seekg(0);
    Image img;
    if (magic[0] == 'P' && (magic[1] == 'F' || magic[1] == 'f')) readPFM(is, img);
    else if (magic[0] == '#' && magic[1] == '?') readRGBE(is, img);
    else error("\"" + path + "\": unsupported image format -- the B200 host reads PFM and Radiance RGBE; OpenEXR, PNG and JPEG need libraries that are unavailable on this box");
    return img;
}

}  // namespace mtsbhost
