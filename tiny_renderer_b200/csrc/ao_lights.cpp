// ao_lights.cpp — deterministic light-direction sets for the many-light ambient-occlusion batch
// (BASELINE.json config 4; semantics in SURVEY.md §8(d)).  Host only.
//
// The reference's occlusion pipeline (/root/reference/src/scene/shader.rs:806-963) takes ONE light direction; the
// batch mode renders it once per direction and sums the gray frames.  Random directions can hit inputs on which the
// reference panics or indexes its shadow buffer on the wrong row, so the generator rejects them up front:
//   (a) light parallel to `up`: the basis of default_prepare (shader.rs:193-195) degenerates;
//   (b) light == -z: Rotation3::rotation_between(+z, light) is None (shader.rs:921);
//   (c) any model vertex outside |ndc| <= margin in the light's view: round(x) >= width or idx >= w*h in the shadow
//       lookups (shader.rs:774-778, 909-912, 932-935).  The margin (0.97) covers the 16 AO samples that step 0.02
//       world units off the fragment (shader.rs:919,929).
#include "../../include/tinyr.h"
#include "tr_error.h"

#include <cmath>
#include <cstdint>

namespace {

struct SplitMix64 {
    uint64_t state;
    uint64_t next() {
        uint64_t z = (state += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }   // [0, 1)
};

struct D3 { double x, y, z; };
D3 sub(D3 a, D3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
D3 cross(D3 a, D3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
D3 scale(D3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
D3 unit(D3 a) { return scale(a, 1.0 / std::sqrt(dot(a, a))); }

// max(|ndc.x|, |ndc.y|) over the vertices for a camera at `from` looking at `at` (default_prepare's
// projection * model * view in double; the viewport maps ndc [-1,1] onto [0, w-1] x [0, h-1]).
double max_ndc(const float* pos, uint32_t n, D3 from, D3 at, D3 up) {
    D3 nz = unit(sub(from, at));
    D3 ny = unit(sub(up, scale(nz, dot(nz, up))));
    D3 nx = unit(cross(ny, nz));
    double worst = 0.0;
    for (uint32_t i = 0; i < n; ++i) {
        D3 p = sub(D3{pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]}, from);
        double x = dot(nx, p), y = dot(ny, p), z = dot(nz, p);
        double w = 1.0 - z / 5.0;
        if (!(w > 0.0)) return INFINITY;
        worst = std::fmax(worst, std::fmax(std::fabs(x / w), std::fabs(y / w)));
    }
    return worst;
}

}  // namespace

extern "C" int tr_ao_generate_lights(const float* positions_xyz, uint32_t n_positions, const float look_at[3], const float up[3],
                                     uint64_t seed, uint32_t n_lights, float ndc_margin, float* out_dirs_xyz, uint64_t* n_drawn) {
    using namespace tr;
    if ((!positions_xyz && n_positions) || !look_at || !up || (!out_dirs_xyz && n_lights))
        return fail(TR_ERR_INVALID_ARGUMENT, "tr_ao_generate_lights: null argument");
    if (!(ndc_margin > 0.0f)) return fail(TR_ERR_INVALID_ARGUMENT, "tr_ao_generate_lights: margin must be positive");
    clear_error();
    SplitMix64 rng{seed};
    const D3 at{look_at[0], look_at[1], look_at[2]}, upv{up[0], up[1], up[2]};
    const double up_len = std::sqrt(dot(upv, upv));
    uint64_t drawn = 0;
    uint32_t accepted = 0;
    const uint64_t limit = 1000ull * n_lights + 1000ull;
    while (accepted < n_lights) {
        if (drawn >= limit) return fail(TR_ERR_REFERENCE_UNDEFINED, "tr_ao_generate_lights: %u of %u directions accepted after %llu draws (model does not fit the light views)",
                                        accepted, n_lights, (unsigned long long)drawn);
        ++drawn;
        const double u1 = rng.uniform(), u2 = rng.uniform();
        const double zc = 1.0 - 2.0 * u1;
        const double r = std::sqrt(std::fmax(0.0, 1.0 - zc * zc));
        const double phi = 2.0 * M_PI * u2;
        const float d[3] = {(float)(r * std::cos(phi)), (float)(r * std::sin(phi)), (float)zc};
        const D3 dd{d[0], d[1], d[2]};
        const double len = std::sqrt(dot(dd, dd));
        if (!(len > 0.0)) continue;
        if (std::fabs(dot(dd, upv)) / (len * up_len) > 0.999) continue;                 // (a)
        if (d[0] == 0.0f && d[1] == 0.0f && d[2] < 0.0f) continue;                       // (b)
        if (max_ndc(positions_xyz, n_positions, dd, at, upv) > (double)ndc_margin) continue;   // (c)
        out_dirs_xyz[3 * accepted] = d[0]; out_dirs_xyz[3 * accepted + 1] = d[1]; out_dirs_xyz[3 * accepted + 2] = d[2];
        ++accepted;
    }
    if (n_drawn) *n_drawn = drawn;
    return TR_OK;
}
