/* systems_ref.c — see systems_ref.h. TEST INFRASTRUCTURE ONLY. */
#include "systems_ref.h"

#include <math.h>

#include "fimo_worlds_b200.h"

void fimo_ref_integrate(float *pos, const float *vel, float dt, size_t n) {
    for (size_t k = 0; k < 3; k++)
        for (size_t e = 0; e < n; e++) {
            const float step = vel[k * n + e] * dt;
            pos[k * n + e] = pos[k * n + e] + step;
        }
}

/* local = T(pos) * R(rot) * S(scale), 3x4 row-major */
static void local_trs(float L[12], float px, float py, float pz, float x, float y, float z, float w, float s) {
    const float xx = x * x, yy = y * y, zz = z * z;
    const float xy = x * y, xz = x * z, yz = y * z;
    const float wx = w * x, wy = w * y, wz = w * z;
    const float r00 = 1.0f - 2.0f * (yy + zz), r01 = 2.0f * (xy - wz), r02 = 2.0f * (xz + wy);
    const float r10 = 2.0f * (xy + wz), r11 = 1.0f - 2.0f * (xx + zz), r12 = 2.0f * (yz - wx);
    const float r20 = 2.0f * (xz - wy), r21 = 2.0f * (yz + wx), r22 = 1.0f - 2.0f * (xx + yy);
    L[0] = r00 * s, L[1] = r01 * s, L[2] = r02 * s, L[3] = px;
    L[4] = r10 * s, L[5] = r11 * s, L[6] = r12 * s, L[7] = py;
    L[8] = r20 * s, L[9] = r21 * s, L[10] = r22 * s, L[11] = pz;
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
                for (int col = 0; col < 3; col++) W[r * 4 + col] = (a * L[col] + b * L[4 + col]) + c * L[8 + col];
                W[r * 4 + 3] = ((a * L[3] + b * L[7]) + c * L[11]) + t;
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
            wc[r] = ((W[r * 4 + 0] * c0 + W[r * 4 + 1] * c1) + W[r * 4 + 2] * c2) + W[r * 4 + 3];
            wh[r] = (fabsf(W[r * 4 + 0]) * h0 + fabsf(W[r * 4 + 1]) * h1) + fabsf(W[r * 4 + 2]) * h2;
        }
        int visible = 1;
        for (int p = 0; p < 6; p++) {
            const float nx = frustum[p * 4 + 0], ny = frustum[p * 4 + 1], nz = frustum[p * 4 + 2], d = frustum[p * 4 + 3];
            const float dist = ((nx * wc[0] + ny * wc[1]) + nz * wc[2]) + d;
            const float rad = (fabsf(nx) * wh[0] + fabsf(ny) * wh[1]) + fabsf(nz) * wh[2];
            if (!(dist + rad >= 0.0f)) visible = 0;
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
