/* systems_ref.h — scalar CPU definition of the synthetic systems (TEST INFRASTRUCTURE ONLY).
 *
 * The reference contains no entity / transform / cull arithmetic at all (SURVEY.md §8 a14:
 * grep for transform|aabb|frustum|cull|entity over /root/reference -> 0 hits), so this file is
 * not a restatement of reference code: it is the DEFINITION of the three build-defined systems
 * named in BASELINE.json (integrate -> transform propagate -> AABB frustum cull), written as
 * plain scalar loops.  "Parity unpinned" for this arithmetic: nothing in the reference can pin it.
 * The CUDA kernels must reproduce these results bit for bit (every float op is a single
 * IEEE-754 binary32 operation rounded to nearest-even — an add, a multiply, or an explicitly written
 * fused multiply-add fmaf() — in exactly the order written; build with -ffp-contract=off so the
 * compiler fuses nothing on its own, device with --fmad=false and __fmaf_rn where fmaf is written).
 *
 * Data layout of one world: component planes, plane k of a C-component
 * resource with E elements is at words [k*E, (k+1)*E).
 */
#ifndef FIMO_ORACLE_SYSTEMS_REF_H
#define FIMO_ORACLE_SYSTEMS_REF_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* integrate   X: pos[3][E]   S: vel[3][E], dt */
void fimo_ref_integrate(float *pos, const float *vel, float dt, size_t n);

/* propagate   X: world[12][E] (3x4 row-major affine)   S: pos, rot[4][E] (x,y,z,w), scale[E], parent[E].
 * Entities are sorted by hierarchy level, so parent[e] < e or parent[e] == 0xFFFFFFFF (root).
 * Returns 0, or -1 if a parent index is not smaller than its child. */
int fimo_ref_propagate(float *world, const float *pos, const float *rot, const float *scale, const uint32_t *parent,
                       size_t n);

/* cull   X: vis_mask[ceil(E/32)], vis_list[E], vis_count[1], stats[8 x u64]
 *        S: world, aabb_c[3][E], aabb_h[3][E], frustum[6][4] (nx,ny,nz,d; inside = n.x + d >= 0) */
void fimo_ref_cull(uint32_t *vis_mask, uint32_t *vis_list, uint32_t *vis_count, uint64_t *stats, const float *world,
                   const float *aabb_c, const float *aabb_h, const float *frustum, size_t n);

/* ---- the 8- and 16-system DAGs (BASELINE.json configs[2], configs[4]); same rules: one rounding per written
 * operation, fmaf exactly where written, sqrtf and / are the IEEE correctly rounded ones ---- */

/* forces   X: vel[3][E]   S: gravity[3], dt        vel = fmaf(g, dt, vel) */
void fimo_ref_forces(float *vel, const float *gravity, float dt, size_t n);
/* vel_clamp   X: vel   S: vmax                     v = fminf(fmaxf(v, -vmax), vmax) */
void fimo_ref_vel_clamp(float *vel, float vmax, size_t n);
/* angvel_damp   X: angvel[3][E]   S: damp          w = w * damp */
void fimo_ref_angvel_damp(float *angvel, float damp, size_t n);
/* integrate_ang   X: rot[4][E] (x,y,z,w)   S: angvel, dt      q += (dt/2) (omega (x) q); q /= |q| */
void fimo_ref_integrate_ang(float *rot, const float *angvel, float dt, size_t n);
/* world_aabb   X: waabb[6][E] (centre xyz, half extent xyz)   S: world, aabb_c, aabb_h */
void fimo_ref_world_aabb(float *waabb, const float *world, const float *aabb_c, const float *aabb_h, size_t n);
/* cull_waabb   X: vis_mask, vis_list, vis_count, stats   S: waabb, frustum   (the plane test of fimo_ref_cull) */
void fimo_ref_cull_waabb(uint32_t *vis_mask, uint32_t *vis_list, uint32_t *vis_count, uint64_t *stats, const float *waabb,
                         const float *frustum, size_t n);
/* bounds_reduce   X: wbounds[6] (min xyz, max xyz)   S: waabb; vis_mask == NULL: every entity, else the visible ones */
void fimo_ref_bounds(float *bounds, const float *waabb, const uint32_t *vis_mask, size_t n);
/* lod   X: lod[E] u32   S: waabb, campos[3], lod_dist2[3]     lod = #{i : |c' - cam|^2 >= lod_dist2[i]} */
void fimo_ref_lod(uint32_t *lod, const float *waabb, const float *campos, const float *lod_dist2, size_t n);
/* sleep_flags   X: sleep[ceil(E/32)]   S: vel, sleep_eps      bit = |vx| + |vy| + |vz| < eps */
void fimo_ref_sleep_flags(uint32_t *sleep, const float *vel, float eps, size_t n);

#ifdef __cplusplus
}
#endif
#endif
