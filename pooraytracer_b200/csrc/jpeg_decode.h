/*
 * jpeg_decode.h — baseline / progressive JPEG to the texel values stb_image produces (jpeg_decode.cpp).
 */
#ifndef PRT_JPEG_DECODE_H
#define PRT_JPEG_DECODE_H
#include <stddef.h>
#include <stdint.h>

#include <string>
#include <vector>

namespace prt {
/* channels: 1 (grey file) or 3, as stbi_load(path, &w, &h, &channels, 0) reports them. Returns false with a reason. */
bool decode_jpeg(const uint8_t* bytes, size_t n, int& width, int& height, int& channels, std::vector<uint8_t>& pixels, std::string& why);
}  // namespace prt
#endif
