// ptb200_png.cpp — PNG ingestion for Image("*.png") without libpng.
//
// The reference decodes PNGs through libpng 1.2 (src/png_decoder.cpp:40-121) with these transforms:
// 16-bit samples stripped to their high byte (png_set_strip_16), sub-byte samples unpacked, palette
// expanded to RGB, and — when the file has no alpha channel — a filler byte of 0 AFTER each pixel
// (png_set_filler(png_ptr, 0, PNG_FILLER_AFTER)), so RGB files load with alpha 0.  Image then maps every
// byte to float as (int)b / 255.0f (src/image.cpp:60-73).  Only zlib (inflate, crc32) is used here.
//
// Not accepted (ImageLoadError): grayscale files — the reference never asks libpng for gray -> RGB, so it
// reads 2-byte pixels into 4-byte rows and the result is not an image — and interlaced files (no
// png_set_interlace_handling there either).  The reference's own PNG inputs (sky01/*.png, test.png)
// are 8-bit RGB / RGBA, non-interlaced.
#include "ptb200_scene.hpp"

#include <zlib.h>

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <vector>

namespace PathTrace
{
namespace
{
uint32_t be32(const unsigned char *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3]; }

// Undoes the per-scanline filters (PNG specification, "Filter Algorithms"), in place.
void unfilter(std::vector<unsigned char> &raw, size_t rowBytes, size_t rows, size_t bpp)
{
    std::vector<unsigned char> zero(rowBytes, 0);
    for(size_t y = 0; y < rows; y++)
    {
        unsigned char *line = &raw[y * (rowBytes + 1)];
        const int type = line[0];
        unsigned char *cur = line + 1;
        const unsigned char *up = y ? line - rowBytes : zero.data();
        for(size_t i = 0; i < rowBytes; i++)
        {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = up[i], c = i >= bpp ? up[i - bpp] : 0;
            int predicted;
            switch(type)
            {
            case 0:
                predicted = 0;
                break;
            case 1:
                predicted = a;
                break;
            case 2:
                predicted = b;
                break;
            case 3:
                predicted = (a + b) >> 1;
                break;
            case 4:
            {
                const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
                predicted = pa <= pb && pa <= pc ? a : pb <= pc ? b : c;
                break;
            }
            default:
                throw ImageLoadError("png: invalid filter type");
            }
            cur[i] = (unsigned char)(cur[i] + predicted);
        }
    }
}
} // namespace

std::shared_ptr<ptb200::Image_data> loadPng(const std::string &fileName)
{
    std::ifstream is(fileName.c_str(), std::ios::binary);
    if(!is)
        throw ImageLoadError("can't open file : \"" + fileName + "\"");
    std::vector<unsigned char> file((std::istreambuf_iterator<char>(is)), std::istreambuf_iterator<char>());
    static const unsigned char signature[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if(file.size() < 8 || memcmp(file.data(), signature, 8) != 0)
        throw ImageLoadError("png: not a PNG file");
    uint32_t w = 0, h = 0;
    int depth = 0, colorType = -1;
    bool gotHeader = false, gotEnd = false;
    std::vector<unsigned char> palette, compressed;
    for(size_t at = 8; !gotEnd;)
    {
        if(at + 12 > file.size())
            throw ImageLoadError("png: truncated file");
        const uint32_t length = be32(&file[at]);
        if(length > file.size() || at + 12 + length > file.size())
            throw ImageLoadError("png: truncated chunk");
        const unsigned char *type = &file[at + 4], *body = &file[at + 8];
        if(be32(body + length) != (uint32_t)crc32(crc32(0L, Z_NULL, 0), type, (uInt)(4 + length)))
            throw ImageLoadError("png: CRC error");
        if(!gotHeader && memcmp(type, "IHDR", 4) != 0)
            throw ImageLoadError("png: missing IHDR");
        if(memcmp(type, "IHDR", 4) == 0)
        {
            if(length != 13 || gotHeader)
                throw ImageLoadError("png: invalid IHDR");
            gotHeader = true;
            w = be32(body), h = be32(body + 4);
            depth = body[8], colorType = body[9];
            if(w == 0 || h == 0 || w > (1u << 20) || h > (1u << 20))
                throw ImageLoadError("png: invalid image size");
            if(body[10] != 0 || body[11] != 0)
                throw ImageLoadError("png: unknown compression or filter method");
            if(body[12] != 0)
                throw ImageLoadError("png: interlaced files are not supported");
            const bool ok = ((colorType == 2 || colorType == 6) && (depth == 8 || depth == 16)) || (colorType == 3 && (depth == 1 || depth == 2 || depth == 4 || depth == 8));
            if(colorType == 0 || colorType == 4)
                throw ImageLoadError("png: grayscale files are not supported");
            if(!ok)
                throw ImageLoadError("png: invalid color type / bit depth");
        }
        else if(memcmp(type, "PLTE", 4) == 0)
        {
            if(length % 3 != 0 || length > 768)
                throw ImageLoadError("png: invalid palette");
            palette.assign(body, body + length);
        }
        else if(memcmp(type, "IDAT", 4) == 0)
            compressed.insert(compressed.end(), body, body + length);
        else if(memcmp(type, "IEND", 4) == 0)
            gotEnd = true;
        else if(!(type[0] & 0x20))
            throw ImageLoadError("png: unknown critical chunk");
        at += 12 + (size_t)length;
    }
    const size_t channels = colorType == 2 ? 3 : colorType == 6 ? 4 : 1;
    const size_t rowBytes = ((size_t)w * channels * depth + 7) / 8, bpp = std::max<size_t>(1, channels * depth / 8);
    std::vector<unsigned char> raw((rowBytes + 1) * (size_t)h);
    {
        z_stream z;
        memset(&z, 0, sizeof(z));
        if(inflateInit(&z) != Z_OK)
            throw ImageLoadError("png: can't start inflate");
        z.next_in = compressed.data(), z.avail_in = (uInt)compressed.size();
        z.next_out = raw.data(), z.avail_out = (uInt)raw.size();
        const int rc = inflate(&z, Z_FINISH);
        const bool complete = rc == Z_STREAM_END && z.avail_out == 0;
        inflateEnd(&z);
        if(!complete)
            throw ImageLoadError("png: image data is damaged or has the wrong size");
    }
    unfilter(raw, rowBytes, h, bpp);
    if(colorType == 3 && palette.empty())
        throw ImageLoadError("png: palette image without PLTE");
    std::shared_ptr<ptb200::Image_data> img(new ptb200::Image_data(w, h));
    for(size_t y = 0; y < h; y++)
    {
        const unsigned char *line = &raw[y * (rowBytes + 1) + 1];
        float *out = &img->rgba[4 * (size_t)w * y];
        for(size_t x = 0; x < w; x++)
        {
            unsigned char px[4] = {0, 0, 0, 0}; // alpha: the 0 filler unless the file has an alpha channel
            if(colorType == 3)
            {
                const size_t bit = x * depth;
                const unsigned index = (line[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1u);
                if(3 * (size_t)index + 2 >= palette.size())
                    throw ImageLoadError("png: palette index out of range");
                px[0] = palette[3 * index], px[1] = palette[3 * index + 1], px[2] = palette[3 * index + 2];
            }
            else
                for(size_t c = 0; c < channels; c++)
                    px[c] = line[(x * channels + c) * (depth / 8)]; // 16 bit: the high byte (big endian)
            for(int c = 0; c < 4; c++)
                out[4 * x + c] = (int)px[c] / 255.0f; // src/image.cpp:70
        }
    }
    return img;
}
} // namespace PathTrace
