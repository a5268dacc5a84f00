// imageio.cpp -- image readers for the `envmap` emitter and the `bitmap` texture: Portable Float Map, Radiance RGBE
// (.hdr / .rgbe) and PNG.
// Reference: Bitmap::readPFM (src/libcore/bitmap.cpp:3764-3814), Bitmap::readRGBE (:3590-3690, RGBE_ToFloat :3522-3530),
// Bitmap::readPNG (:2450-2580; restated on zlib).  OpenEXR / JPEG need libraries that are not on this box; asking for
// them is reported as an error.
#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iterator>

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
    img.luminance = !color;
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

// ---- PNG (Bitmap::readPNG, bitmap.cpp:2450-2580, restated on zlib: libpng is not on this box) ------------------------
// Palette and 1/2/4-bit grey images are expanded, 16-bit samples kept, alpha dropped (a texture reads luminance or RGB,
// bitmap.cpp:268-279).  Gamma: sRGB chunk or nothing -> sRGB (bitmap.cpp:2534-2541), gAMA -> 1/gAMA; samples are
// linearised as FormatConverter does (fmtconv.cpp:151-157, 1092-1101: value/max, then the sRGB curve or pow in Float).
inline uint32_t be32(const unsigned char *p) { return ((uint32_t) p[0] << 24) | ((uint32_t) p[1] << 16) | ((uint32_t) p[2] << 8) | p[3]; }

void readPNG(std::istream &is, Image &img, float gammaOverride) {
    std::vector<unsigned char> file((std::istreambuf_iterator<char>(is)), std::istreambuf_iterator<char>());
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (file.size() < 8 || std::memcmp(file.data(), sig, 8) != 0) error("readPNG(): invalid signature!");
    int depth = 0, ctype = 0, interlace = 0;
    float gamma = -1.0f;
    std::vector<unsigned char> idat, palette;
    for (size_t pos = 8; pos + 12 <= file.size();) {
        const uint32_t len = be32(&file[pos]);
        const std::string tag((const char *) &file[pos + 4], 4);
        const unsigned char *d = &file[pos + 8];
        if (pos + 12 + (size_t) len > file.size()) error("readPNG(): truncated chunk!");
        if (tag == "IHDR") {
            if (len < 13) error("readPNG(): bad IHDR!");
            img.width = (int) be32(d); img.height = (int) be32(d + 4); depth = d[8]; ctype = d[9]; interlace = d[12];
        } else if (tag == "PLTE") palette.assign(d, d + len);
        else if (tag == "IDAT") idat.insert(idat.end(), d, d + len);
        else if (tag == "sRGB") gamma = -1.0f;
        else if (tag == "gAMA" && len == 4 && gamma != -2.0f) { const double g = be32(d) / 100000.0; if (g > 0) gamma = (float) 1 / (float) g; }
        else if (tag == "IEND") break;
        pos += 12 + (size_t) len;
    }
    // an sRGB chunk wins over gAMA (png_get_sRGB is asked first)
    for (size_t pos = 8; pos + 12 <= file.size(); pos += 12 + (size_t) be32(&file[pos]))
        if (std::memcmp(&file[pos + 4], "sRGB", 4) == 0) gamma = -1.0f;
    if (gammaOverride != 0) gamma = gammaOverride;            // Bitmap::setGamma (textures/bitmap.cpp:251-252): -1 = sRGB, else the exponent
    if (img.width <= 0 || img.height <= 0) error("readPNG(): invalid image size!");
    if (interlace != 0) error("readPNG(): interlaced PNG files are not supported by the B200 host");
    const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!channels || !(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)) error("readPNG(): unsupported color type / bit depth!");
    if (ctype == 3 && palette.empty()) error("readPNG(): palette missing!");
    const size_t bpp = std::max<size_t>(1, (size_t) channels * depth / 8), stride = ((size_t) img.width * channels * depth + 7) / 8;
    std::vector<unsigned char> raw((stride + 1) * (size_t) img.height);
    uLongf outLen = (uLongf) raw.size();
    if (uncompress(raw.data(), &outLen, idat.data(), (uLong) idat.size()) != Z_OK || outLen != raw.size()) error("readPNG(): could not inflate the image data!");
    // undo the scanline filters (PNG specification, section 9)
    std::vector<unsigned char> prev(stride, 0);
    for (int y = 0; y < img.height; ++y) {
        unsigned char *row = &raw[(stride + 1) * (size_t) y];
        const int f = row[0];
        unsigned char *cur = row + 1;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            int pred = 0;
            switch (f) {
                case 0: break;
                case 1: pred = a; break;
                case 2: pred = b; break;
                case 3: pred = (a + b) >> 1; break;
                case 4: { const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c); pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
                default: error("readPNG(): bad filter type!");
            }
            cur[i] = (unsigned char) (cur[i] + pred);
        }
        std::memcpy(prev.data(), cur, stride);
    }
    const bool grey = ctype == 0 || ctype == 4;
    img.luminance = grey;
    const int maxValue = (ctype == 3 || depth < 8) ? 255 : (1 << depth) - 1;
    std::vector<float> lut((size_t) maxValue + 1);
    const float mult = 1.0f / (float) maxValue;
    for (int i = 0; i <= maxValue; ++i) {
        const float v = (float) i * mult;
        lut[i] = gamma == -1.0f ? (v <= 0.04045f ? v * (float) (1.0 / 12.92) : std::pow((float) ((v + 0.055f) * (float) (1.0 / 1.055)), 2.4f)) : std::pow(v, gamma);
    }
    img.rgb.resize((size_t) img.width * img.height * 3);
    for (int y = 0; y < img.height; ++y) {
        const unsigned char *cur = &raw[(stride + 1) * (size_t) y + 1];
        for (int x = 0; x < img.width; ++x) {
            int v[4] = {0, 0, 0, 0};
            for (int c = 0; c < channels; ++c) {
                const size_t s = (size_t) x * channels + c;
                if (depth == 16) v[c] = (cur[2 * s] << 8) | cur[2 * s + 1];
                else if (depth == 8) v[c] = cur[s];
                else { const size_t bit = s * depth; v[c] = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1); }
            }
            float *o = &img.rgb[3 * ((size_t) y * img.width + x)];
            if (ctype == 3) {
                if ((size_t) v[0] * 3 + 2 >= palette.size()) error("readPNG(): palette index out of range!");
                for (int c = 0; c < 3; ++c) o[c] = lut[palette[(size_t) v[0] * 3 + c]];
            } else if (grey) {
                const int g = depth < 8 ? v[0] * 255 / ((1 << depth) - 1) : v[0];      // png_set_expand_gray_1_2_4_to_8
                o[0] = o[1] = o[2] = lut[g];
            } else {
                for (int c = 0; c < 3; ++c) o[c] = lut[v[c]];
            }
        }
    }
}

}  // namespace

// FormatConverter::undoGamma (fmtconv.cpp:1091-1100): what a `gamma` override does to an image whose file format is linear
static float undoGamma(float value, float gamma) {
    if (gamma == -1.0f) return value <= 0.04045f ? value * (float) (1.0 / 12.92) : std::pow((float) ((value + 0.055f) * (float) (1.0 / 1.055)), 2.4f);
    return std::pow(value, gamma);
}

// gammaOverride: the `gamma` property of the bitmap texture / envmap (0 = keep the file's own gamma, -1 = sRGB, otherwise
// the exponent that linearises the stored values; textures/bitmap.cpp:218,251-252, envmap.cpp:127-128)
Image readImage(const std::string &path, const char *what, float gammaOverride) {
    std::ifstream is(path, std::ios::binary);
    if (!is) error(std::string(what) + " file \"" + path + "\" could not be found!");          // envmap.cpp:121-122, bitmap.cpp:196-197
    char magic[2] = {0, 0};
    is.read(magic, 2);
    is.seekg(0);
    Image img;
    const bool linearFormat = (magic[0] == 'P' && (magic[1] == 'F' || magic[1] == 'f')) || (magic[0] == '#' && magic[1] == '?');
    if (magic[0] == 'P' && (magic[1] == 'F' || magic[1] == 'f')) readPFM(is, img);
    else if (magic[0] == '#' && magic[1] == '?') readRGBE(is, img);
    else if ((unsigned char) magic[0] == 0x89 && magic[1] == 'P') readPNG(is, img, gammaOverride);
    else error("\"" + path + "\": unsupported image format -- the B200 host reads PFM, Radiance RGBE and PNG; OpenEXR and JPEG need libraries that are unavailable on this box");
    if (linearFormat && gammaOverride != 0 && gammaOverride != 1.0f)      // PFM / RGBE carry gamma 1; an override re-interprets the values
        for (float &v : img.rgb) v = undoGamma(v, gammaOverride);
    return img;
}

}  // namespace mtsbhost
