// Minimal 8-bit RGB PNG encoder (zlib deflate + CRC), replacing the stbi_write_png call at
// /root/reference/raytracing.cu:395. Parity is on decoded pixels, not on PNG bytes.
#include <zlib.h>

#include <cstdint>
#include <cstdio>
#include <vector>

#include "../../../include/rt_scene.h"

namespace {

void put_u32(std::vector<uint8_t> &out, uint32_t v)
{
    out.push_back(uint8_t(v >> 24));
    out.push_back(uint8_t(v >> 16));
    out.push_back(uint8_t(v >> 8));
    out.push_back(uint8_t(v));
}

void put_chunk(std::vector<uint8_t> &out, const char tag[4], const uint8_t *data, size_t n)
{
    put_u32(out, (uint32_t)n);
    size_t start = out.size();
    out.insert(out.end(), tag, tag + 4);
    if (n) out.insert(out.end(), data, data + n);
    put_u32(out, (uint32_t)crc32(0L, out.data() + start, (uInt)(n + 4)));
}

}  // namespace

extern "C" int rt_write_png(const char *path, int width, int height, const uint8_t *rgb8)
{
    if (!path || !rgb8 || width <= 0 || height <= 0) return RT_ERR_INVALID;
    const size_t row = (size_t)width * 3;
    std::vector<uint8_t> raw((row + 1) * (size_t)height);
    for (int y = 0; y < height; y++) {
        raw[(row + 1) * y] = 0;  // filter: none
        std::copy(rgb8 + row * y, rgb8 + row * (y + 1), raw.begin() + (row + 1) * y + 1);
    }
    uLongf packed_size = compressBound((uLong)raw.size());
    std::vector<uint8_t> packed(packed_size);
    if (compress2(packed.data(), &packed_size, raw.data(), (uLong)raw.size(), 6) != Z_OK) return RT_ERR_IO;

    std::vector<uint8_t> png = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    uint8_t header[13];
    std::vector<uint8_t> tmp;
    put_u32(tmp, (uint32_t)width);
    put_u32(tmp, (uint32_t)height);
    std::copy(tmp.begin(), tmp.end(), header);
    header[8] = 8;   // bit depth
    header[9] = 2;   // colour type RGB
    header[10] = header[11] = header[12] = 0;
    put_chunk(png, "IHDR", header, 13);
    put_chunk(png, "IDAT", packed.data(), packed_size);
    put_chunk(png, "IEND", nullptr, 0);

    FILE *f = std::fopen(path, "wb");
    if (!f) return RT_ERR_IO;
    size_t written = std::fwrite(png.data(), 1, png.size(), f);
    std::fclose(f);
    return written == png.size() ? RT_OK : RT_ERR_IO;
}
