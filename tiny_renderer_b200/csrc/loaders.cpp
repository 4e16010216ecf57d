// loaders.cpp — host-side asset I/O kept from the reference's boundary: OBJ parse, TGA decode, TGA write.
// Replaces obj::raw::parse_obj (obj-rs 0.7.0) and image::open(..).into_rgb8() (image 0.24.5) as called by
// /root/reference/src/app.rs:87-131; adds the headless TGA writer that stands in for the window
// (/root/reference/src/app.rs:148-153, 216-218 — no graphics stack on this box).
#include "../../include/tinyr.h"
#include "tr_error.h"

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <charconv>
#include <string>
#include <vector>

namespace tr {

static thread_local std::string g_error;

int fail(int status, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_error = buf;
    return status;
}
void clear_error() { g_error.clear(); }
const char* last_error() { return g_error.c_str(); }

static bool read_file(const char* path, std::vector<uint8_t>* out) {
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

// ---- OBJ ---------------------------------------------------------------------------------------------------

static bool parse_f32(const std::string& s, float* v) {
    if (s.empty()) return false;
    // std::from_chars: correctly rounded like Rust's str::parse::<f32>, and independent of the host application's
    // locale (strtof reads the decimal point from setlocale).  Rust accepts one leading '+', from_chars does not.
    const char* b = s.data();
    const char* e = b + s.size();
    if (*b == '+' && e - b > 1 && b[1] != '-' && b[1] != '+') ++b;
    const std::from_chars_result r = std::from_chars(b, e, *v, std::chars_format::general);
    return r.ec == std::errc() && r.ptr == e;
}

// One face corner: "p", "p/t", "p//n" or "p/t/n"; returns a bitmask 1 (p) | 2 (t) | 4 (n), 0 on error.
static int parse_corner(const std::string& s, long idx[3]) {
    int mask = 0;
    size_t start = 0;
    for (int k = 0; k < 3; ++k) {
        size_t slash = s.find('/', start);
        std::string part = s.substr(start, slash == std::string::npos ? std::string::npos : slash - start);
        if (!part.empty()) {
            char* end = nullptr;
            idx[k] = strtol(part.c_str(), &end, 10);
            if (!end || *end != '\0') return 0;
            mask |= 1 << k;
        }
        if (slash == std::string::npos) break;
        start = slash + 1;
    }
    return (mask & 1) ? mask : 0;
}

static bool resolve_index(long raw, size_t count, uint32_t* out) {
    long v = raw > 0 ? raw - 1 : (long)count + raw;   // 1-based, or negative = relative to the end
    if (raw == 0 || v < 0 || (size_t)v >= count) return false;
    *out = (uint32_t)v;
    return true;
}

}  // namespace tr

using namespace tr;

extern "C" int tr_load_obj(const char* path, tr_mesh* out) {
    if (!path || !out) return fail(TR_ERR_INVALID_ARGUMENT, "tr_load_obj: null argument");
    memset(out, 0, sizeof *out);
    std::vector<uint8_t> bytes;
    if (!read_file(path, &bytes)) return fail(TR_ERR_IO, "cannot read %s: %s", path, strerror(errno));
    std::vector<float> pos, uv, nrm;
    std::vector<uint32_t> tri;
    std::string text(bytes.begin(), bytes.end());
    size_t line_no = 0, at = 0;
    std::string line;
    while (at <= text.size()) {
        // logical line: physical lines joined on a trailing backslash
        line.clear();
        while (at <= text.size()) {
            size_t nl = text.find('\n', at);
            std::string phys = text.substr(at, nl == std::string::npos ? std::string::npos : nl - at);
            at = nl == std::string::npos ? text.size() + 1 : nl + 1;
            ++line_no;
            if (!phys.empty() && phys.back() == '\r') phys.pop_back();
            if (!phys.empty() && phys.back() == '\\') { phys.pop_back(); line += phys; line += ' '; continue; }
            line += phys;
            break;
        }
        size_t hash = line.find('#');
        if (hash != std::string::npos) line.resize(hash);
        std::vector<std::string> tok;
        size_t i = 0;
        while (i < line.size()) {
            while (i < line.size() && isspace((unsigned char)line[i])) ++i;
            size_t j = i;
            while (j < line.size() && !isspace((unsigned char)line[j])) ++j;
            if (j > i) tok.push_back(line.substr(i, j - i));
            i = j;
        }
        if (tok.empty()) continue;
        const std::string& kw = tok[0];
        size_t n_args = tok.size() - 1;
        if (kw == "v" || kw == "vt" || kw == "vn") {
            float f[4] = {0, 0, 0, 0};
            size_t lo = kw == "v" ? 3 : (kw == "vt" ? 1 : 3), hi = kw == "v" ? 4 : 3;
            if (n_args < lo || n_args > hi) return fail(TR_ERR_FORMAT, "%s:%zu: `%s` takes %zu..%zu numbers", path, line_no, kw.c_str(), lo, hi);
            for (size_t k = 0; k < n_args; ++k)
                if (!parse_f32(tok[1 + k], &f[k])) return fail(TR_ERR_FORMAT, "%s:%zu: bad number `%s`", path, line_no, tok[1 + k].c_str());
            if (kw == "v") pos.insert(pos.end(), f, f + 3);          // w ignored by the reference (util.rs:25-31)
            else if (kw == "vt") uv.insert(uv.end(), f, f + 2);      // 3rd component ignored (shader.rs:144)
            else nrm.insert(nrm.end(), f, f + 3);
        } else if (kw == "f") {
            if (n_args < 3) return fail(TR_ERR_FORMAT, "%s:%zu: face with fewer than 3 corners", path, line_no);
            uint32_t nine[9];
            for (size_t c = 0; c < n_args; ++c) {
                long idx[3] = {0, 0, 0};
                int mask = parse_corner(tok[1 + c], idx);
                if (mask == 0) return fail(TR_ERR_FORMAT, "%s:%zu: bad face corner `%s`", path, line_no, tok[1 + c].c_str());
                // The reference only accepts Polygon::PTN (scene.rs:216-219).
                if (mask != 7) return fail(TR_ERR_BAD_POLYGON, "%s:%zu: polygon is not position/texture/normal (scene.rs:218 panics)", path, line_no);
                uint32_t p, t, n;
                if (!resolve_index(idx[0], pos.size() / 3, &p) || !resolve_index(idx[1], uv.size() / 2, &t) ||
                    !resolve_index(idx[2], nrm.size() / 3, &n))
                    return fail(TR_ERR_FORMAT, "%s:%zu: index out of range in `%s`", path, line_no, tok[1 + c].c_str());
                if (c < 3) { nine[3 * c] = p; nine[3 * c + 1] = t; nine[3 * c + 2] = n; }   // first 3 corners (scene.rs:224-226)
            }
            tri.insert(tri.end(), nine, nine + 9);
        }
        // every other statement (g, o, s, usemtl, mtllib, l, p, ...) carries nothing the render path reads
    }
    auto dup_f = [](const std::vector<float>& v) { float* p = (float*)malloc(v.size() * 4 + 4); if (p) memcpy(p, v.data(), v.size() * 4); return p; };
    out->positions = dup_f(pos); out->n_positions = (uint32_t)(pos.size() / 3);
    out->uvs = dup_f(uv); out->n_uvs = (uint32_t)(uv.size() / 2);
    out->normals = dup_f(nrm); out->n_normals = (uint32_t)(nrm.size() / 3);
    out->tri_idx = (uint32_t*)malloc(tri.size() * 4 + 4);
    if (out->tri_idx) memcpy(out->tri_idx, tri.data(), tri.size() * 4);
    out->n_triangles = (uint32_t)(tri.size() / 9);
    if (!out->positions || !out->uvs || !out->normals || !out->tri_idx) { tr_mesh_free(out); return fail(TR_ERR_INVALID_ARGUMENT, "out of host memory"); }
    return TR_OK;
}

extern "C" void tr_mesh_free(tr_mesh* m) {
    if (!m) return;
    free(m->positions); free(m->uvs); free(m->normals); free(m->tri_idx);
    memset(m, 0, sizeof *m);
}
