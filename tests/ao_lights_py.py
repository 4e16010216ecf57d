"""Independent Python twin of tr_ao_generate_lights (tiny_renderer_b200/csrc/ao_lights.cpp): SplitMix64 -> uniform
sphere directions in double -> f32, with the rejection rules of SURVEY.md §8(d) config 4.  Uses math.sin/cos/sqrt
(the C library's, like the C++ side) so the accepted list is identical bit for bit."""
import math

import numpy as np

MASK = (1 << 64) - 1


def splitmix64(state):
    state = (state + 0x9E3779B97F4A7C15) & MASK
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
    return state, z ^ (z >> 31)


def max_ndc(P, d, at, up):
    nz = (d - at) / np.linalg.norm(d - at)
    ny = up - nz * np.dot(nz, up)
    ny /= np.linalg.norm(ny)
    nx = np.cross(ny, nz)
    nx /= np.linalg.norm(nx)
    q = P - d
    w = 1.0 - (q @ nz) / 5.0
    if not (w > 0).all():
        return math.inf
    return max(np.abs((q @ nx) / w).max(), np.abs((q @ ny) / w).max())


def generate(positions, n_lights, seed=0x5EED0A0, look_at=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0), margin=0.97, slack=0.0):
    """Returns (dirs (n,3) f32, n_drawn).  `slack` widens the ndc test by a relative amount (the twin sums in a
    different order than the C++ loop, so a direction within 1e-12 of the margin could flip; none does here)."""
    P = np.asarray(positions, np.float64)
    at, upv = np.asarray(look_at, np.float64), np.asarray(up, np.float64)
    out, drawn, state = [], 0, seed
    while len(out) < n_lights:
        drawn += 1
        state, a = splitmix64(state)
        state, b = splitmix64(state)
        u1, u2 = (a >> 11) / 9007199254740992.0, (b >> 11) / 9007199254740992.0
        zc = 1.0 - 2.0 * u1
        r = math.sqrt(max(0.0, 1.0 - zc * zc))
        phi = 2.0 * math.pi * u2
        d = np.array([r * math.cos(phi), r * math.sin(phi), zc]).astype(np.float32)
        dd = d.astype(np.float64)
        ln = np.linalg.norm(dd)
        if not ln > 0:
            continue
        if abs(np.dot(dd, upv)) / (ln * np.linalg.norm(upv)) > 0.999:
            continue
        if d[0] == 0 and d[1] == 0 and d[2] < 0:
            continue
        if max_ndc(P, dd, at, upv) > margin * (1.0 + slack):
            continue
        out.append(d)
    return np.array(out, np.float32).reshape(-1, 3), drawn
