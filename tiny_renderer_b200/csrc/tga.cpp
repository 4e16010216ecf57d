// tga.cpp — Truevision TGA decode (what image 0.24.5's `image::open(path)?.into_rgb8()` yields for the
// reference's assets, /root/reference/src/app.rs:99,107,118,126) and the headless TGA writer.
// Decoder behaviour restated from the image crate's tga codec: types 1/2/3 raw and 9/10/11 RLE; 8-bit gray,
// 24-bit BGR, 32-bit BGRA, colour-mapped with 24/32-bit entries; RLE packets may cross scanlines; images whose
// descriptor bit 5 is clear (bottom-left origin) are flipped to top-down; bit 4 is ignored; into_rgb8 drops
// alpha and replicates gray.
#include "../../include/tinyr.h"
#include "tr_error.h"

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace tr;

namespace {

bool slurp(const char* path, std::vector<uint8_t>* out) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out->resize(n > 0 ? (size_t)n : 0);
    size_t got = n > 0 ? fread(out->data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == out->size();
}

inline uint32_t le16(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }

}  // namespace

extern "C" int tr_load_tga(const char* path, tr_pixels* out) {
    if (!path || !out) return fail(TR_ERR_INVALID_ARGUMENT, "tr_load_tga: null argument");
    memset(out, 0, sizeof *out);
    std::vector<uint8_t> file;
    if (!slurp(path, &file)) return fail(TR_ERR_IO, "cannot read %s: %s", path, strerror(errno));
    if (file.size() < 18) return fail(TR_ERR_FORMAT, "%s: shorter than a TGA header", path);
    const uint8_t* h = file.data();
    const uint32_t id_len = h[0], map_type = h[1], img_type = h[2];
    const uint32_t map_len = le16(h + 5), map_entry_bits = h[7];
    const uint32_t width = le16(h + 12), height = le16(h + 14), depth = h[16], desc = h[17];
    const bool rle = img_type == 9 || img_type == 10 || img_type == 11;
    const uint32_t base_type = rle ? img_type - 8 : img_type;
    if (base_type < 1 || base_type > 3) return fail(TR_ERR_FORMAT, "%s: unsupported TGA image type %u", path, img_type);
    if (width == 0 || height == 0) return fail(TR_ERR_FORMAT, "%s: empty image", path);
    const uint32_t bpp = depth / 8;          // bytes per stored pixel (index for colour-mapped)
    if (depth % 8 != 0 || bpp < 1 || bpp > 4) return fail(TR_ERR_FORMAT, "%s: unsupported pixel depth %u", path, depth);
    if (base_type == 3 && bpp != 1) return fail(TR_ERR_FORMAT, "%s: gray image with depth %u", path, depth);
    if (base_type == 2 && bpp != 3 && bpp != 4) return fail(TR_ERR_FORMAT, "%s: true-colour image with depth %u", path, depth);
    size_t at = 18 + id_len;
    const uint8_t* cmap = nullptr;
    uint32_t cmap_bpp = 0;
    if (map_type == 1) {
        cmap_bpp = (map_entry_bits + 7) / 8;
        if (at + (size_t)map_len * cmap_bpp > file.size()) return fail(TR_ERR_FORMAT, "%s: truncated colour map", path);
        cmap = file.data() + at;
        at += (size_t)map_len * cmap_bpp;
    }
    if (base_type == 1 && (!cmap || (cmap_bpp != 3 && cmap_bpp != 4) || bpp > 2))
        return fail(TR_ERR_FORMAT, "%s: unsupported colour-mapped layout", path);

    // 1. raw stored pixels, file order
    const size_t n_px = (size_t)width * height;
    std::vector<uint8_t> stored(n_px * bpp);
    if (!rle) {
        if (at + stored.size() > file.size()) return fail(TR_ERR_FORMAT, "%s: truncated pixel data", path);
        memcpy(stored.data(), file.data() + at, stored.size());
    } else {
        size_t o = 0;
        while (o < stored.size()) {
            if (at >= file.size()) return fail(TR_ERR_FORMAT, "%s: truncated RLE stream", path);
            uint8_t head = file[at++];
            size_t count = (size_t)(head & 0x7f) + 1;
            if (head & 0x80) {
                if (at + bpp > file.size()) return fail(TR_ERR_FORMAT, "%s: truncated RLE run", path);
                for (size_t k = 0; k < count && o < stored.size(); ++k, o += bpp) memcpy(&stored[o], &file[at], bpp);
                at += bpp;
            } else {
                size_t bytes = count * bpp;
                if (at + bytes > file.size()) return fail(TR_ERR_FORMAT, "%s: truncated RLE literal", path);
                size_t take = bytes < stored.size() - o ? bytes : stored.size() - o;
                memcpy(&stored[o], &file[at], take);
                o += take;
                at += bytes;
            }
        }
    }

    // 2. to RGB8, top-down
    uint8_t* rgb = (uint8_t*)malloc(n_px * 3);
    if (!rgb) return fail(TR_ERR_INVALID_ARGUMENT, "out of host memory");
    const bool top_origin = (desc & 0x20) != 0;
    for (uint32_t y = 0; y < height; ++y) {
        uint32_t dst_y = top_origin ? y : height - 1 - y;
        for (uint32_t x = 0; x < width; ++x) {
            const uint8_t* s = &stored[((size_t)y * width + x) * bpp];
            uint8_t* d = rgb + ((size_t)dst_y * width + x) * 3;
            if (base_type == 3) { d[0] = d[1] = d[2] = s[0]; }
            else if (base_type == 2) { d[0] = s[2]; d[1] = s[1]; d[2] = s[0]; }
            else {
                uint32_t idx = bpp == 1 ? s[0] : le16(s);
                uint32_t first = le16(h + 3);
                idx = idx >= first ? idx - first : 0;
                if (idx >= map_len) { free(rgb); return fail(TR_ERR_FORMAT, "%s: colour index out of range", path); }
                const uint8_t* e = cmap + (size_t)idx * cmap_bpp;
                d[0] = e[2]; d[1] = e[1]; d[2] = e[0];
            }
        }
    }
    out->rgb8_topdown = rgb; out->width = width; out->height = height;
    return TR_OK;
}

extern "C" void tr_pixels_free(tr_pixels* p) {
    if (!p) return;
    free(p->rgb8_topdown);
    memset(p, 0, sizeof *p);
}

// 24-bit true-colour, bottom-left origin like the reference's own assets; optional RLE (type 10), packets
// confined to one scanline.
extern "C" int tr_write_tga(const char* path, const uint8_t* rgb8_topdown, uint32_t width, uint32_t height, int rle) {
    if (!path || !rgb8_topdown || width == 0 || height == 0 || width > 65535 || height > 65535)
        return fail(TR_ERR_INVALID_ARGUMENT, "tr_write_tga: bad argument");
    std::vector<uint8_t> out;
    out.reserve(18 + (size_t)width * height * 3 + 26);
    uint8_t hdr[18] = {0};
    hdr[2] = rle ? 10 : 2;
    hdr[12] = (uint8_t)(width & 0xff); hdr[13] = (uint8_t)(width >> 8);
    hdr[14] = (uint8_t)(height & 0xff); hdr[15] = (uint8_t)(height >> 8);
    hdr[16] = 24; hdr[17] = 0;
    out.insert(out.end(), hdr, hdr + 18);
    std::vector<uint8_t> row((size_t)width * 3);
    for (uint32_t y = 0; y < height; ++y) {
        const uint8_t* src = rgb8_topdown + (size_t)(height - 1 - y) * width * 3;   // file row 0 = bottom
        for (uint32_t x = 0; x < width; ++x) { row[3 * x] = src[3 * x + 2]; row[3 * x + 1] = src[3 * x + 1]; row[3 * x + 2] = src[3 * x]; }
        if (!rle) { out.insert(out.end(), row.begin(), row.end()); continue; }
        uint32_t x = 0;
        while (x < width) {
            uint32_t run = 1;
            while (x + run < width && run < 128 && memcmp(&row[3 * x], &row[3 * (x + run)], 3) == 0) ++run;
            if (run >= 2) {
                out.push_back((uint8_t)(0x80 | (run - 1)));
                out.insert(out.end(), &row[3 * x], &row[3 * x] + 3);
                x += run;
            } else {
                uint32_t lit = 1;
                while (x + lit < width && lit < 128 &&
                       !(x + lit + 1 < width && memcmp(&row[3 * (x + lit)], &row[3 * (x + lit + 1)], 3) == 0)) ++lit;
                out.push_back((uint8_t)(lit - 1));
                out.insert(out.end(), &row[3 * x], &row[3 * x] + (size_t)lit * 3);
                x += lit;
            }
        }
    }
    static const char footer[26] = "\0\0\0\0\0\0\0\0TRUEVISION-XFILE.";
    out.insert(out.end(), footer, footer + 26);
    FILE* f = fopen(path, "wb");
    if (!f) return fail(TR_ERR_IO, "cannot open %s for writing: %s", path, strerror(errno));
    size_t n = fwrite(out.data(), 1, out.size(), f);
    if (fclose(f) != 0 || n != out.size()) return fail(TR_ERR_IO, "short write to %s", path);
    return TR_OK;
}
