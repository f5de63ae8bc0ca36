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

/* ---------------------------------------------------------------------------------------------
 * the 8- and 16-system DAGs
 * ------------------------------------------------------------------------------------------- */
void fimo_ref_forces(float *vel, const float *gravity, float dt, size_t n) {
    for (size_t k = 0; k < 3; k++)
        for (size_t e = 0; e < n; e++) vel[k * n + e] = fmaf(gravity[k], dt, vel[k * n + e]);
}

void fimo_ref_vel_clamp(float *vel, float vmax, size_t n) {
    for (size_t i = 0; i < 3 * n; i++) vel[i] = fminf(fmaxf(vel[i], -vmax), vmax);
}

void fimo_ref_angvel_damp(float *angvel, float damp, size_t n) {
    for (size_t i = 0; i < 3 * n; i++) angvel[i] = angvel[i] * damp;
}

void fimo_ref_integrate_ang(float *rot, const float *angvel, float dt, size_t n) {
    const float h = 0.5f * dt;
    for (size_t e = 0; e < n; e++) {
        float x = rot[e], y = rot[n + e], z = rot[2 * n + e], w = rot[3 * n + e];
        const float ax = angvel[e], ay = angvel[n + e], az = angvel[2 * n + e];
        /* (ax, ay, az, 0) (x) (x, y, z, w) */
        const float dx = fmaf(ay, z, fmaf(-az, y, ax * w));
        const float dy = fmaf(az, x, fmaf(-ax, z, ay * w));
        const float dz = fmaf(ax, y, fmaf(-ay, x, az * w));
        const float dw = -fmaf(az, z, fmaf(ay, y, ax * x));
        x = fmaf(dx, h, x), y = fmaf(dy, h, y), z = fmaf(dz, h, z), w = fmaf(dw, h, w);
        const float len = sqrtf(fmaf(w, w, fmaf(z, z, fmaf(y, y, x * x))));
        rot[e] = x / len, rot[n + e] = y / len, rot[2 * n + e] = z / len, rot[3 * n + e] = w / len;
    }
}

void fimo_ref_world_aabb(float *waabb, const float *world, const float *aabb_c, const float *aabb_h, size_t n) {
    for (size_t e = 0; e < n; e++) {
        float W[12];
        for (int i = 0; i < 12; i++) W[i] = world[(size_t)i * n + e];
        const float c0 = aabb_c[e], c1 = aabb_c[n + e], c2 = aabb_c[2 * n + e];
        const float h0 = aabb_h[e], h1 = aabb_h[n + e], h2 = aabb_h[2 * n + e];
        for (int r = 0; r < 3; r++) {
            waabb[(size_t)r * n + e] = fmaf(W[r * 4 + 2], c2, fmaf(W[r * 4 + 1], c1, fmaf(W[r * 4 + 0], c0, W[r * 4 + 3])));
            waabb[(size_t)(3 + r) * n + e] =
                    fmaf(fabsf(W[r * 4 + 2]), h2, fmaf(fabsf(W[r * 4 + 1]), h1, fabsf(W[r * 4 + 0]) * h0));
        }
    }
}

void fimo_ref_cull_waabb(uint32_t *vis_mask, uint32_t *vis_list, uint32_t *vis_count, uint64_t *stats, const float *waabb,
                         const float *frustum, size_t n) {
    const size_t words = (n + 31) / 32;
    for (size_t i = 0; i < words; i++) vis_mask[i] = 0;
    uint32_t count = 0;
    for (size_t e = 0; e < n; e++) {
        float wc[3], wh[3];
        for (int r = 0; r < 3; r++) wc[r] = waabb[(size_t)r * n + e], wh[r] = waabb[(size_t)(3 + r) * n + e];
        int visible = 1;
        for (int p = 0; p < 6; p++) {
            const float nx = frustum[p * 4 + 0], ny = frustum[p * 4 + 1], nz = frustum[p * 4 + 2], d = frustum[p * 4 + 3];
            float sd = fmaf(nz, wc[2], fmaf(ny, wc[1], fmaf(nx, wc[0], d)));
            sd = fmaf(fabsf(nz), wh[2], fmaf(fabsf(ny), wh[1], fmaf(fabsf(nx), wh[0], sd)));
            if (!(sd >= 0.0f)) visible = 0;
        }
        if (visible) {
            vis_mask[e >> 5] |= 1u << (e & 31);
            vis_list[count++] = (uint32_t)e;
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

void fimo_ref_bounds(float *bounds, const float *waabb, const uint32_t *vis_mask, size_t n) {
    float lo[3] = {FIMO_B200_BOUNDS_EMPTY_MIN, FIMO_B200_BOUNDS_EMPTY_MIN, FIMO_B200_BOUNDS_EMPTY_MIN};
    float hi[3] = {-FIMO_B200_BOUNDS_EMPTY_MIN, -FIMO_B200_BOUNDS_EMPTY_MIN, -FIMO_B200_BOUNDS_EMPTY_MIN};
    for (size_t e = 0; e < n; e++) {
        if (vis_mask && !((vis_mask[e >> 5] >> (e & 31)) & 1u)) continue;
        for (int r = 0; r < 3; r++) {
            const float c = waabb[(size_t)r * n + e], h = waabb[(size_t)(3 + r) * n + e];
            lo[r] = fminf(lo[r], c - h);
            hi[r] = fmaxf(hi[r], c + h);
        }
    }
    for (int r = 0; r < 3; r++) bounds[r] = lo[r], bounds[3 + r] = hi[r];
}

void fimo_ref_lod(uint32_t *lod, const float *waabb, const float *campos, const float *lod_dist2, size_t n) {
    for (size_t e = 0; e < n; e++) {
        const float dx = waabb[e] - campos[0], dy = waabb[n + e] - campos[1], dz = waabb[2 * n + e] - campos[2];
        const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
        lod[e] = (uint32_t)(d2 >= lod_dist2[0]) + (uint32_t)(d2 >= lod_dist2[1]) + (uint32_t)(d2 >= lod_dist2[2]);
    }
}

void fimo_ref_sleep_flags(uint32_t *sleep, const float *vel, float eps, size_t n) {
    const size_t words = (n + 31) / 32;
    for (size_t i = 0; i < words; i++) sleep[i] = 0;
    for (size_t e = 0; e < n; e++) {
        const float a = (fabsf(vel[e]) + fabsf(vel[n + e])) + fabsf(vel[2 * n + e]);
        if (a < eps) sleep[e >> 5] |= 1u << (e & 31);
    }
}
