/* systems_ref.c — see systems_ref.h. TEST INFRASTRUCTURE ONLY. */
#include "systems_ref.h"

#include <math.h>

#include "fimo_worlds_b200.h"

void fimo_ref_integrate(float *pos, const float *vel, float dt, size_t n) {
    for (size_t k = 0; k < 3; k++)
        for (size_t e = 0; e < n; e++) pos[k * n + e] = fmaf(vel[k * n + e], dt, pos[k * n + e]);
}

/* local = T(pos) * R(rot) * S(scale), 3x4 row-major.  33 roundings, in this order. */
static void local_trs(float L[12], float px, float py, float pz, float x, float y, float z, float w, float s) {
    const float x2 = x + x, y2 = y + y, z2 = z + z;
    const float xx = x * x2, yy = y * y2, zz = z * z2;
    const float xy = x * y2, xz = x * z2, yz = y * z2;
    const float wx = w * x2, wy = w * y2, wz = w * z2;
    L[0] = fmaf(-(yy + zz), s, s), L[1] = (xy - wz) * s, L[2] = (xz + wy) * s, L[3] = px;
    L[4] = (xy + wz) * s, L[5] = fmaf(-(xx + zz), s, s), L[6] = (yz - wx) * s, L[7] = py;
    L[8] = (xz - wy) * s, L[9] = (yz + wx) * s, L[10] = fmaf(-(xx + yy), s, s), L[11] = pz;
}

int fimo_ref_propagate(float *world, const float *pos, const float *rot, const float *scale, const uint32_t *parent,
                       size_t n) {
    for (size_t e = 0; e < n; e++) {
        float L[12], W[12];
        local_trs(L, pos[e], pos[n + e], pos[2 * n + e], rot[e], rot[n + e], rot[2 * n + e], rot[3 * n + e], scale[e]);
        const uint32_t p = parent[e];
        if (p == FIMO_B200_PARENT_ROOT) {
            for (int i = 0; i < 12; i++) W[i] = L[i];
        } else {
            if ((size_t)p >= e) return -1;
            float P[12];
            for (int i = 0; i < 12; i++) P[i] = world[(size_t)i * n + p];
            for (int r = 0; r < 3; r++) {
                const float a = P[r * 4 + 0], b = P[r * 4 + 1], c = P[r * 4 + 2], t = P[r * 4 + 3];
                for (int col = 0; col < 3; col++) W[r * 4 + col] = fmaf(c, L[8 + col], fmaf(b, L[4 + col], a * L[col]));
                W[r * 4 + 3] = fmaf(c, L[11], fmaf(b, L[7], fmaf(a, L[3], t)));
            }
        }
        for (int i = 0; i < 12; i++) world[(size_t)i * n + e] = W[i];
    }
    return 0;
}

void fimo_ref_cull(uint32_t *vis_mask, uint32_t *vis_list, uint32_t *vis_count, uint64_t *stats, const float *world,
                   const float *aabb_c, const float *aabb_h, const float *frustum, size_t n) {
    const size_t words = (n + 31) / 32;
    for (size_t i = 0; i < words; i++) vis_mask[i] = 0;
    uint32_t count = 0;
    for (size_t e = 0; e < n; e++) {
        float W[12];
        for (int i = 0; i < 12; i++) W[i] = world[(size_t)i * n + e];
        const float c0 = aabb_c[e], c1 = aabb_c[n + e], c2 = aabb_c[2 * n + e];
        const float h0 = aabb_h[e], h1 = aabb_h[n + e], h2 = aabb_h[2 * n + e];
        float wc[3], wh[3];
        for (int r = 0; r < 3; r++) {
            wc[r] = fmaf(W[r * 4 + 2], c2, fmaf(W[r * 4 + 1], c1, fmaf(W[r * 4 + 0], c0, W[r * 4 + 3])));
            wh[r] = fmaf(fabsf(W[r * 4 + 2]), h2, fmaf(fabsf(W[r * 4 + 1]), h1, fabsf(W[r * 4 + 0]) * h0));
        }
        int visible = 1;
        for (int p = 0; p < 6; p++) {
            const float nx = frustum[p * 4 + 0], ny = frustum[p * 4 + 1], nz = frustum[p * 4 + 2], d = frustum[p * 4 + 3];
            /* signed distance of the centre plus the projected radius, accumulated in one FMA chain */
            float sd = fmaf(nz, wc[2], fmaf(ny, wc[1], fmaf(nx, wc[0], d)));
            sd = fmaf(fabsf(nz), wh[2], fmaf(fabsf(ny), wh[1], fmaf(fabsf(nx), wh[0], sd)));
            if (!(sd >= 0.0f)) visible = 0;
        }
        if (visible) {
            vis_mask[e >> 5] |= 1u << (e & 31);
            vis_list[count++] = (uint32_t)e; /* ascending entity index; entries >= count are left untouched */
        }
    }
    *vis_count = count;
    uint64_t mask_sum = 0;
    for (size_t i = 0; i < words; i++) mask_sum += vis_mask[i];
    stats[FIMO_B200_STAT_VISIBLE] = count;
    stats[FIMO_B200_STAT_ENTITIES] = n;
    stats[FIMO_B200_STAT_MASK_SUM] = mask_sum;
    stats[FIMO_B200_STAT_WORLDS] = 1;
    stats[FIMO_B200_STAT_STEPS] += 1;
}
