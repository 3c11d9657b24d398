// ptb200_image.cpp — host image I/O either side of the render path:
// Radiance RGBE (.hdr/.pic) reading into the shared RGBA32F Image and the RLE writer behind
// MutableImage::writeHDR.  The value mapping follows the reference so files round-trip to the
// same floats: radiance = mantissa * 179 * 2^(e-136) (reference src/image.cpp:306-314) and the
// writer is its inverse with the same run-length layout (src/image.cpp:398-481), so an image
// written here is byte-identical to one written by the reference from the same floats.
// Said plainly: writeHDR / encodeChannel RESTATE the decisions of src/image.cpp:398-481 one for one (the
// exponent as log(max)/log(2), the +1e-5 guard, runs of three or more, the 0x7F / 0x80 run bytes) — byte identity
// with the reference's files leaves no freedom there; this is a restatement of one function, not a redesign.
#include "ptb200_scene.hpp"

#include <ctype.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <fstream>

namespace
{
struct FileReader
{
    std::vector<unsigned char> bytes;
    size_t pos;
    explicit FileReader(const std::string &name) : pos(0)
    {
        std::ifstream is(name.c_str(), std::ios::binary);
        if(!is)
            throw ImageLoadError("can't open " + name);
        bytes.assign(std::istreambuf_iterator<char>(is), std::istreambuf_iterator<char>());
    }
    int get()
    {
        if(pos >= bytes.size())
            throw ImageLoadError("unexpected EOF");
        return bytes[pos++];
    }
    std::string line()
    {
        std::string s;
        for(int c = get(); c != '\n'; c = get())
            s += (char)c;
        return s;
    }
};

std::string trimmed(const std::string &s)
{
    size_t a = 0, b = s.size();
    while(a < b && isspace((unsigned char)s[a]))
        a++;
    while(b > a && isspace((unsigned char)s[b - 1]))
        b--;
    return s.substr(a, b - a);
}

// One scanline of w RGBE quads.  New-style RLE (marker 2,2,hi,lo; four planar runs) is what the
// reference reads (src/image.cpp:226-268); flat scanlines are accepted too.
void readScanline(FileReader &in, unsigned char *quads, int w)
{
    unsigned char head[4];
    for(int i = 0; i < 4; i++)
        head[i] = (unsigned char)in.get();
    bool rle = head[0] == 2 && head[1] == 2 && !(head[2] & 0x80);
    if(!rle)
    {
        memcpy(quads, head, 4);
        for(int i = 4; i < 4 * w; i++)
            quads[i] = (unsigned char)in.get();
        return;
    }
    if(((head[2] << 8) | head[3]) != w)
        throw ImageLoadError("invalid line length in new compressed line");
    for(int channel = 0; channel < 4; channel++)
    {
        int x = 0;
        while(x < w)
        {
            int code = in.get();
            if(code > 0x80)
            {
                int value = in.get();
                for(int n = code - 0x80; n > 0; n--)
                {
                    if(x >= w)
                        throw ImageLoadError("line too long");
                    quads[4 * x++ + channel] = (unsigned char)value;
                }
            }
            else
            {
                for(int n = code; n > 0; n--)
                {
                    if(x >= w)
                        throw ImageLoadError("line too long");
                    quads[4 * x++ + channel] = (unsigned char)in.get();
                }
            }
        }
    }
}

std::shared_ptr<ptb200::Image_data> loadRadiance(const std::string &fileName)
{
    FileReader in(fileName);
    if(in.line() != "#?RADIANCE")
        throw ImageLoadError("magic string doesn't match");
    float scale[3] = {1, 1, 1};
    bool gotFormat = false;
    std::string resolution;
    for(;;)
    {
        std::string l = trimmed(in.line());
        if(l.empty() || l[0] == '#')
            continue;
        if(l[0] == '-' || l[0] == '+')
        {
            resolution = l;
            break;
        }
        size_t eq = l.find('=');
        if(eq == std::string::npos)
            throw ImageLoadError("unexpected token");
        std::string key = trimmed(l.substr(0, eq)), value = trimmed(l.substr(eq + 1));
        if(key == "FORMAT")
        {
            if(gotFormat)
                throw ImageLoadError("format already specified");
            gotFormat = true;
            if(value != "32-bit_rle_rgbe")
                throw ImageLoadError("invalid format specifier");
        }
        else if(key == "EXPOSURE")
        { // src/image.cpp:140-154: radiance is divided by every EXPOSURE value
            float e;
            if(sscanf(value.c_str(), "%f", &e) != 1)
                throw ImageLoadError("can't read exposure value");
            for(int c = 0; c < 3; c++)
                scale[c] /= e;
        }
        else if(key == "COLORCORR")
        { // src/image.cpp:156-171
            float cc[3];
            if(sscanf(value.c_str(), "%f %f %f", &cc[0], &cc[1], &cc[2]) != 3)
                throw ImageLoadError("can't read color correction values");
            for(int c = 0; c < 3; c++)
                scale[c] /= cc[c];
        }
    }
    int w = 0, h = 0;
    if(sscanf(resolution.c_str(), "-Y %d +X %d", &h, &w) != 2 || h <= 0 || w <= 0 || w >= 1 << 15)
        throw ImageLoadError("invalid resolution string");
    std::shared_ptr<ptb200::Image_data> img = std::make_shared<ptb200::Image_data>(w, h);
    std::vector<unsigned char> quads((size_t)4 * w);
    for(int y = 0; y < h; y++)
    {
        readScanline(in, quads.data(), w);
        float *row = &img->rgba[(size_t)4 * w * y];
        for(int x = 0; x < w; x++)
        {
            int exponent = quads[4 * x + 3] - 128;
            float factor = 179.0f * (float)pow(2.0, exponent - 8);
            row[4 * x + 0] = quads[4 * x + 0] * factor * scale[0];
            row[4 * x + 1] = quads[4 * x + 1] * factor * scale[1];
            row[4 * x + 2] = quads[4 * x + 2] * factor * scale[2];
            row[4 * x + 3] = 1;
        }
    }
    return img;
}

// Planar run-length encoding of one channel of a scanline, same layout decisions as
// src/image.cpp:433-474: runs need >= 3 equal bytes and are at most 127 long; shorter repeats
// are folded into the pending literal block, which is flushed at 128 bytes.
void encodeChannel(std::ofstream &os, const unsigned char *quads, size_t w, size_t channel)
{
    struct At
    {
        const unsigned char *q;
        size_t c;
        unsigned char operator()(size_t x) const { return q[4 * x + c]; }
    } at = {quads, channel};
    size_t x = 0;
    while(x < w)
    {
        size_t literal = 0, run = 0;
        while(run < 0x7F && literal <= 0x80 && x + literal + run < w)
        {
            while(run < 0x7F && x + literal + run < w && at(x + literal) == at(x + literal + run))
                run++;
            if(run >= 3)
                break;
            literal += run;
            run = 0;
        }
        if(run == 0 && literal > 0x80)
            literal = 0x80;
        if(literal > 0)
        {
            os.put((char)(unsigned char)literal);
            for(size_t i = 0; i < literal; i++)
                os.put((char)at(x + i));
            x += literal;
        }
        if(run > 0)
        {
            os.put((char)(unsigned char)(run + 0x80));
            os.put((char)at(x));
            x += run;
        }
    }
}
} // namespace

namespace PathTrace
{
std::shared_ptr<ptb200::Image_data> loadPng(const std::string &fileName); // ptb200_png.cpp
}

Image::Image(std::string fileName, std::string format)
{
    if(format == "")
    {
        size_t dot = fileName.find_last_of('.');
        if(dot == std::string::npos)
            throw ImageLoadError("can't determine format");
        format = fileName.substr(dot + 1);
        for(size_t i = 0; i < format.size(); i++)
            format[i] = (char)tolower(format[i]);
    }
    if(format == "hdr" || format == "pic")
        data = loadRadiance(fileName);
    else if(format == "png")
        data = PathTrace::loadPng(fileName);
    else
        throw ImageLoadError("invalid format");
}

void MutableImage::writeHDR(std::string fileName) const
{
    std::ofstream os(fileName.c_str(), std::ios::binary);
    if(!os)
        throw ImageStoreError("can't open file for writing");
    os << "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y " << h << " +X " << w << "\n";
    std::vector<unsigned char> quads((size_t)4 * w);
    for(size_t y = 0; y < h; y++)
    {
        const float *row = &px[(size_t)4 * w * y];
        for(size_t x = 0; x < w; x++)
        {
            const float *p = row + 4 * x;
            unsigned char *q = &quads[4 * x];
            float top = std::max(p[0], std::max(p[1], p[2])) / 179.0f;
            if(top < 1e-30)
            {
                q[0] = q[1] = q[2] = q[3] = 0;
                continue;
            }
            int lg = (int)ceil((double)logf(top) / log(2.0) + 1e-5); // src/image.cpp:423 (std::log(float) / log(2))
            float scale = (float)(pow(0.5, (double)(lg - 8)) / 179.0f);
            for(int c = 0; c < 3; c++)
                q[c] = (unsigned char)std::max(0, std::min(0xFF, (int)floorf(p[c] * scale)));
            q[3] = (unsigned char)(lg + 128);
        }
        const unsigned char marker[4] = {2, 2, (unsigned char)(w >> 8), (unsigned char)(w & 0xFF)};
        os.write((const char *)marker, 4);
        for(size_t channel = 0; channel < 4; channel++)
            encodeChannel(os, quads.data(), w, channel);
        if(!os)
            throw ImageStoreError("can't write to file");
    }
}
