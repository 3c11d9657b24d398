// ptb200_tile_text.cpp — the tile wire text of the reference's render farm, host only.
//
// The reference's server (src/test.cpp:678-716) answers a connection with "1\n" (or "0\n" when busy),
// reads " x y size", answers "S\n", then once a second writes every finished pixel it has not written
// yet as "P<x>,<y>=<r>,<g>,<b>\n" (printf %i and %g, RenderBlock::copyToSocket, :372-391) followed by
// "E\n", and "F\n" after the last batch.  The client (NetRenderBlock::run, :596-630) skips 'S', 'E'
// and newlines, parses 'P' records with fscanf("%i,%i=%f,%f,%f"), rejects pixels outside its tile and
// stops at 'F'.  These functions produce and consume exactly that text for tiles rendered through
// ptb200_render* / ptb200_job_read; sockets stay with the caller.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "ptb200.h"

extern "C" int ptb200_tile_request_parse(const char *line, int screen_width, int screen_height, int *x, int *y, int *size)
{
    int tx, ty, ts; // src/test.cpp:688: 3 != fscanf(f2, " %i %i %i", ...) and the range checks that follow
    if(!line || !x || !y || !size || 3 != sscanf(line, " %i %i %i", &tx, &ty, &ts) || tx < 0 || ty < 0 || ts <= 0 || ts > 1000000 || tx > screen_width || ty > screen_height ||
       ts > screen_width || ts > screen_height)
        return PTB_EINVAL;
    *x = tx, *y = ty, *size = ts;
    return PTB_OK;
}

extern "C" int ptb200_tile_encode(const float *rgba, int width, int height, int x, int y, int size, uint8_t *sent, int final, char *out, size_t capacity, size_t *length)
{
    if(!rgba || width <= 0 || height <= 0 || size <= 0 || !length)
        return PTB_EINVAL;
    std::string s;
    std::vector<uint8_t *> marked; // wrotePixel flags of the pixels in the text: committed only once the text is in `out`
    char buf[128];
    for(int py = y; py < y + size; py++) // RenderBlock::copyToSocket, src/test.cpp:374-388
        for(int px = x; px < x + size; px++)
        {
            if(px < 0 || py < 0 || px >= width || py >= height)
                continue; // beyond the screen: never valid in a frame buffer
            const float *p = rgba + 4 * ((size_t)py * width + px);
            uint8_t *flag = sent ? sent + (size_t)(py - y) * size + (px - x) : NULL;
            if(!(p[3] > 0.f) || (flag && *flag))
                continue; // validPixel && !wrotePixel
            if(flag)
                marked.push_back(flag);
            snprintf(buf, sizeof(buf), "P%i,%i=%g,%g,%g\n", px, py, p[0], p[1], p[2]);
            s += buf;
        }
    s += "E\n";
    if(final)
        s += "F\n"; // src/test.cpp:715
    *length = s.size();
    if(!out || !capacity)
        return PTB_OK; // measuring only: sent[] is untouched, the next call sends the same pixels
    if(capacity < s.size() + 1)
        return PTB_ENOMEM; // nothing was sent
    memcpy(out, s.c_str(), s.size() + 1);
    for(uint8_t *flag : marked)
        *flag = 1;
    return PTB_OK;
}

extern "C" int ptb200_tile_decode(const char *text, size_t length, int x, int y, int size, float *rgb_tile, uint8_t *valid, int *finished, size_t *consumed)
{
    if(!text || size <= 0 || !rgb_tile || !valid || !finished)
        return PTB_EINVAL;
    const std::string s(text, length); // NUL-terminated copy for sscanf
    size_t i = 0;
    *finished = 0;
    int rc = PTB_OK;
    while(i < s.size())
    {
        const char c = s[i];
        if(c == 'S' || c == 'E' || c == '\n')
        { // src/test.cpp:600-604
            i++;
            continue;
        }
        if(c == 'F')
        { // :626-628
            i++;
            *finished = 1;
            break;
        }
        if(c != 'P')
        { // :629-631
            rc = PTB_EINVAL;
            break;
        }
        int px, py, n = 0;
        float r, g, b;
        if(s.find('\n', i) == std::string::npos)
            break; // an incomplete record: wait for more bytes
        if(5 != sscanf(s.c_str() + i + 1, "%i,%i=%f,%f,%f%n", &px, &py, &r, &g, &b, &n) || px < x || py < y || px >= x + size || py >= y + size)
        { // :610-614
            rc = PTB_EINVAL;
            break;
        }
        const size_t at = (size_t)(py - y) * size + (px - x);
        valid[at] = 1; // :617-620
        rgb_tile[3 * at] = r, rgb_tile[3 * at + 1] = g, rgb_tile[3 * at + 2] = b;
        i += 1 + (size_t)n;
    }
    if(consumed)
        *consumed = i;
    return rc;
}
