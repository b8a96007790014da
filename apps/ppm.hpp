// apps/ppm.hpp -- image output of the drivers (host side, outside the timed path).
//   write_ppm   P6, rows from j = height-1 down to 0, each channel uint8(max(min(p*255,255),0)) (truncation):
//               the semantics of static.cpp:135-147.
//   write_ppm16 P6 with maxval 65535, channel = uint16(p*65535) clamped, big-endian: stands in for the Magick++
//               16-bit quantum frames of anim.cpp:112-119 (ImageMagick is unavailable on this box).
#ifndef CERES_APPS_PPM_HPP
#define CERES_APPS_PPM_HPP

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace ceres_apps {

inline uint8_t to_u8(float p) {
    float x = p * 255.0f;
    x = (255.0f < x) ? 255.0f : x;      // std::min(x, 255)
    x = (x < 0.0f) ? 0.0f : x;          // std::max(x, 0)
    return x == x ? static_cast<uint8_t>(x) : 0;
}

// flip = true: static.cpp:135-147 (rows from j = height-1 down); false: anim.cpp:117-119 (image row = frame row)
inline bool write_ppm(const std::string& path, const float* pixels, size_t width, size_t height, bool flip = true) {
    std::vector<uint8_t> bytes(3 * width * height);
    #pragma omp parallel for
    for (size_t row = 0; row < height; ++row) {
        const float* src = pixels + 3 * width * (flip ? height - 1 - row : row);
        uint8_t* dst = bytes.data() + 3 * width * row;
        for (size_t k = 0; k < 3 * width; ++k) dst[k] = to_u8(src[k]);
    }
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fprintf(f, "P6 %zu %zu 255\n", width, height);
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    return std::fclose(f) == 0 && ok;
}

// P6 payload that is already quantised and flipped (ceres_render_ppm8)
// (bytes_per_channel 1: ceres_render_ppm8, maxval 255; 2: ceres_render_ppm16, maxval 65535, big-endian)
inline bool write_ppm_bytes(const std::string& path, const uint8_t* bytes, size_t width, size_t height, int bytes_per_channel = 1) {
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fprintf(f, "P6 %zu %zu %d\n", width, height, bytes_per_channel == 2 ? 65535 : 255);
    const size_t n = 3 * width * height * (size_t)bytes_per_channel;
    const bool ok = std::fwrite(bytes, 1, n, f) == n;
    return std::fclose(f) == 0 && ok;
}

inline bool write_ppm16(const std::string& path, const float* pixels, size_t width, size_t height, bool flip = true) {
    std::vector<uint8_t> bytes(6 * width * height);
    #pragma omp parallel for
    for (size_t row = 0; row < height; ++row) {
        const float* src = pixels + 3 * width * (flip ? height - 1 - row : row);
        uint8_t* dst = bytes.data() + 6 * width * row;
        for (size_t k = 0; k < 3 * width; ++k) {
            float x = src[k] * 65535.0f;
            x = (65535.0f < x) ? 65535.0f : x;
            x = (x < 0.0f) ? 0.0f : x;
            const uint16_t q = x == x ? static_cast<uint16_t>(x) : 0;
            dst[2 * k] = static_cast<uint8_t>(q >> 8);
            dst[2 * k + 1] = static_cast<uint8_t>(q & 0xFF);
        }
    }
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fprintf(f, "P6 %zu %zu 65535\n", width, height);
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    return std::fclose(f) == 0 && ok;
}

}  // namespace ceres_apps
#endif
