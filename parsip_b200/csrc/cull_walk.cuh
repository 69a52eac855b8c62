// Per-thread pruning walk of the culling pass (see cull.cu): interval bounds of primitives / warps over a
// region box and the zero-folding rewrite of the postfix program. Pure __device__ functions, kept in a header
// so tests/hostemu can compile them for the host.
#ifndef PSGPU_CULL_WALK_CUH
#define PSGPU_CULL_WALK_CUH
#include "interp.cuh"

namespace psgpu {

struct Box3 { float lo[3], hi[3]; };

__device__ __forceinline__ void imul(float alo, float ahi, float blo, float bhi, float& lo, float& hi) {
    const float p0 = alo * blo, p1 = alo * bhi, p2 = ahi * blo, p3 = ahi * bhi;
    lo = fminf(fminf(p0, p1), fminf(p2, p3));
    hi = fmaxf(fmaxf(p0, p1), fmaxf(p2, p3));
}

// range of cos and sin over [t0, t1]
__device__ void trig_range(float t0, float t1, float& cmin, float& cmax, float& smin, float& smax) {
    const float TWO_PI = 6.2831853f, PI = 3.14159265f;
    if (!(t1 - t0 < TWO_PI)) { cmin = smin = -1.f; cmax = smax = 1.f; return; }
    float s0, c0, s1, c1;
    sincosf(t0, &s0, &c0); sincosf(t1, &s1, &c1);
    cmin = fminf(c0, c1); cmax = fmaxf(c0, c1); smin = fminf(s0, s1); smax = fmaxf(s0, s1);
    // extremum at phase + 2*pi*n inside [t0, t1]  <=>  floor((t1-phase)/2pi) > floor((t0-phase)/2pi) or exact hit
    auto has = [&](float phase) { return floorf((t1 - phase) / TWO_PI) != floorf((t0 - phase) / TWO_PI) || fmodf(t0 - phase, TWO_PI) == 0.f; };
    if (has(0.f)) cmax = 1.f;
    if (has(PI)) cmin = -1.f;
    if (has(0.5f * PI)) smax = 1.f;
    if (has(1.5f * PI)) smin = -1.f;
    const float e = 1e-5f;
    cmin = fmaxf(cmin - e, -1.f); cmax = fminf(cmax + e, 1.f); smin = fmaxf(smin - e, -1.f); smax = fminf(smax + e, 1.f);
}

__device__ __forceinline__ void pad_box(Box3& b) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float e = 1e-5f * (fabsf(b.lo[a]) + fabsf(b.hi[a])) + 1e-6f;
        b.lo[a] -= e; b.hi[a] += e;
    }
}

// conservative image of a box under a warp / operator affine (interval arithmetic), SURVEY 7.3 item 4b
__device__ void warp_box(uint32_t opc, const float4 q0, const float4* __restrict__ q, Box3& b) {
    Box3 o = b;
    if (opc == OPC_AFFINE) {
        const float4 r[3] = {q[0], q[1], q[2]};
        float c[3], h[3];
#pragma unroll
        for (int a = 0; a < 3; a++) { c[a] = 0.5f * (b.lo[a] + b.hi[a]); h[a] = 0.5f * (b.hi[a] - b.lo[a]); }
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float cc = r[a].x * c[0] + r[a].y * c[1] + r[a].z * c[2] + r[a].w;
            const float hh = fabsf(r[a].x) * h[0] + fabsf(r[a].y) * h[1] + fabsf(r[a].z) * h[2];
            o.lo[a] = cc - hh; o.hi[a] = cc + hh;
        }
    } else if (opc == OPC_TWIST) {
        const float k = q0.y;
        const int ax = (int)q0.z;
        const int u = (ax == 0) ? 1 : 0, v = (ax == 2) ? 1 : 2;
        const float t0 = fminf(b.lo[ax] * k, b.hi[ax] * k), t1 = fmaxf(b.lo[ax] * k, b.hi[ax] * k);
        float cmin, cmax, smin, smax;
        trig_range(t0, t1, cmin, cmax, smin, smax);
        float a0, a1, b0, b1;
        imul(b.lo[u], b.hi[u], cmin, cmax, a0, a1); imul(b.lo[v], b.hi[v], smin, smax, b0, b1);
        o.lo[u] = a0 - b1; o.hi[u] = a1 - b0;
        imul(b.lo[u], b.hi[u], smin, smax, a0, a1); imul(b.lo[v], b.hi[v], cmin, cmax, b0, b1);
        o.lo[v] = a0 + b0; o.hi[v] = a1 + b1;
    } else if (opc == OPC_TAPER) {
        const float k = q0.y;
        const int al = (int)q0.z, tp = (int)q0.w;
        int tgt;
        if (al == 0) tgt = (tp == 2) ? 2 : 1; else if (al == 1) tgt = (tp == 2) ? 2 : 0; else tgt = (tp == 2) ? 1 : 0;
        const float s0 = 1.0f + fminf(b.lo[al] * k, b.hi[al] * k), s1 = 1.0f + fmaxf(b.lo[al] * k, b.hi[al] * k);
        imul(b.lo[tgt], b.hi[tgt], s0, s1, o.lo[tgt], o.hi[tgt]);
    } else if (opc == OPC_BEND) {
        const float k = q0.y, y0 = q0.z, left = q0.w, right = q[0].x, kd = q[0].y;
        const float yh0 = fminf(fmaxf(b.lo[1], left), right), yh1 = fminf(fmaxf(b.hi[1], left), right);
        const float ta = k * (yh0 - y0), tb = k * (yh1 - y0);
        float cmin, cmax, smin, smax;
        trig_range(fminf(ta, tb), fmaxf(ta, tb), cmin, cmax, smin, smax);
        // the straight extension outside [left, right]: e(y) = y - left below, 0 inside, y - right above - monotone in y, so
        // its range over the box is [e(lo), e(hi)] (not the hull with 0: a box far outside the bent region keeps its size)
        const float dy0 = fminf(b.lo[1] - left, 0.f) + fmaxf(b.lo[1] - right, 0.f), dy1 = fminf(b.hi[1] - left, 0.f) + fmaxf(b.hi[1] - right, 0.f);
        const float z0 = b.lo[2] - kd, z1 = b.hi[2] - kd;
        float a0, a1, b0, b1;
        imul(smin, smax, z0, z1, a0, a1); imul(cmin, cmax, dy0, dy1, b0, b1);
        o.lo[1] = -a1 + y0 + b0; o.hi[1] = -a0 + y0 + b1;
        imul(cmin, cmax, z0, z1, a0, a1); imul(smin, smax, dy0, dy1, b0, b1);
        o.lo[2] = a0 + kd + b0; o.hi[2] = a1 + kd + b1;
    } else {  // OPC_SHEAR
        const float k = q0.y;
        const int al = (int)q0.z, dp = (int)q0.w;
        int tgt, src;
        if (al == 1) { tgt = 1; src = (dp == 2) ? 2 : 0; } else if (al == 2) { tgt = 2; src = (dp == 1) ? 1 : 0; } else { tgt = 0; src = (dp == 2) ? 2 : 1; }
        o.lo[tgt] = b.lo[tgt] + fminf(k * b.lo[src], k * b.hi[src]);
        o.hi[tgt] = b.hi[tgt] + fmaxf(k * b.lo[src], k * b.hi[src]);
    }
    pad_box(o);
    b = o;
}

// Exact-support test against the box itself (the balls of prim_bound overshoot a box's corners by up to (sqrt(3) - 1) half-sides):
// true if no point of the box lies within the support distance (1; R for a quadric point) of the skeleton, by comparing the box,
// grown by the support in every axis (a superset of its Minkowski sum with the support ball), with the skeleton's own extent -
// per axis for finite skeletons, by slab clipping for the unclamped line of the PS_SimdPoly semantics. Primitives without a matrix only.
__device__ __forceinline__ bool prim_outside_box(uint32_t opc, const float4 q0, const float4* __restrict__ q, const Box3& b) {
    float klo[3] = {q0.y, q0.z, q0.w}, khi[3] = {q0.y, q0.z, q0.w};  // extent of the skeleton
    float m = 1.0f;
    switch (opc) {
    case OPC_POINT: break;
    case OPC_QUADRIC: m = q[0].z; break;
    case OPC_SEGMENT: {
        const float4 d = q[0];
        const float e[3] = {d.x, d.y, d.z};
#pragma unroll
        for (int a = 0; a < 3; a++) { klo[a] += fminf(e[a], 0.f); khi[a] += fmaxf(e[a], 0.f); }
    } break;
    case OPC_CUBE: {
        const float side = fabsf(q[0].x);
#pragma unroll
        for (int a = 0; a < 3; a++) { klo[a] -= side; khi[a] += side; }
    } break;
    case OPC_CYLINDER: case OPC_DISC: case OPC_RING: {
        const float4 d = q[0];
        const float n2 = d.x * d.x + d.y * d.y + d.z * d.z;
        if (!(fabsf(n2 - 1.0f) < 1e-3f)) return false;  // the distance laws assume a unit axis / normal: leave others to the ball bound
        const float rad = fabsf(d.w) * 1.001f;
        const float hgt = (opc == OPC_CYLINDER) ? fmaxf(q[1].x, 0.f) * 1.001f : 0.f;
        const float e[3] = {d.x * hgt, d.y * hgt, d.z * hgt};
#pragma unroll
        for (int a = 0; a < 3; a++) { klo[a] += fminf(e[a], 0.f) - rad; khi[a] += fmaxf(e[a], 0.f) + rad; }
    } break;
    case OPC_LINE_INF: {
        const float4 d = q[0];
        const float e[3] = {d.x, d.y, d.z};
        float t0 = -3.0e38f, t1 = 3.0e38f;
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float pad = 1.0f + 1e-4f + 1e-5f * (fabsf(b.lo[a]) + fabsf(b.hi[a]) + fabsf(klo[a]));
            const float lo = b.lo[a] - pad - klo[a], hi = b.hi[a] + pad - klo[a];
            if (fabsf(e[a]) < 1e-12f) { if (lo > 0.f || hi < 0.f) return true; }
            else {
                const float inv = 1.0f / e[a];
                const float u = lo * inv, v = hi * inv;
                const float s = 1e-5f * (fabsf(u) + fabsf(v));
                t0 = fmaxf(t0, fminf(u, v) - s); t1 = fminf(t1, fmaxf(u, v) + s);
            }
        }
        return t0 > t1;
    }
    default: return false;
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float pad = m * 1.00001f + 1e-4f + 1e-5f * (fabsf(b.lo[a]) + fabsf(b.hi[a]) + fabsf(klo[a]) + fabsf(khi[a]));
        if (klo[a] > b.hi[a] + pad || khi[a] < b.lo[a] - pad) return true;
    }
    return false;
}

// [lo, hi] bound of one primitive's field over the ball around the box; returns true if identically zero.
__device__ bool prim_bound(uint32_t hdr, const float4 q0, const float4* __restrict__ q, const Box3& b, float& flo, float& fhi) {
    const uint32_t opc = PS_HDR_OPC(hdr);
    if (opc == OPC_NULLPRIM) { flo = fhi = 0.f; return true; }
    float c[1][3], r;
    {
        const float hx = 0.5f * (b.hi[0] - b.lo[0]), hy = 0.5f * (b.hi[1] - b.lo[1]), hz = 0.5f * (b.hi[2] - b.lo[2]);
        c[0][0] = 0.5f * (b.lo[0] + b.hi[0]); c[0][1] = 0.5f * (b.lo[1] + b.hi[1]); c[0][2] = 0.5f * (b.lo[2] + b.hi[2]);
        r = sqrtf(hx * hx + hy * hy + hz * hz);
    }
    float x[1] = {c[0][0]}, y[1] = {c[0][1]}, z[1] = {c[0][2]};
    const int nq = ps_base_quads(opc);
    if (PS_HDR_XF(hdr)) {
        const float4* m = q + (nq - 1);
        float ox[1], oy[1], oz[1];
        xform_points<1>(m, x, y, z, ox, oy, oz);
        x[0] = ox[0]; y[0] = oy[0]; z[0] = oz[0];
        float fro = 0.f;
#pragma unroll
        for (int a = 0; a < 3; a++) fro += m[a].x * m[a].x + m[a].y * m[a].y + m[a].z * m[a].z;
        r *= sqrtf(fro) * 1.00001f;
    }
    float d2[1];
    prim_d2<1>(opc, q0, q, x, y, z, d2);
    if (!(d2[0] == d2[0])) { flo = 0.f; fhi = 1.f; return false; }  // NaN: keep, unknown
    const float d = sqrtf(d2[0]);
    const float eps = 4e-6f * (fabsf(x[0]) + fabsf(y[0]) + fabsf(z[0]) + 1.0f) + 1e-5f * r + 1e-5f * d;
    const float dmin = fmaxf(d - r - eps, 0.f), dmax = d + r + eps;
    if (opc == OPC_QUADRIC) {  // support radius R, law scale * (1 - d^2/R^2)^2 (scale > 0 is checked by the linearizer)
        const float4 a = q[0];
        if (dmin >= a.z) { flo = fhi = 0.f; return true; }
        const float vhi = fmaxf(1.0f - dmin * dmin * a.x, 0.f), vlo = fmaxf(1.0f - dmax * dmax * a.x, 0.f);
        fhi = a.y * vhi * vhi * 1.00001f + 1e-7f;
        flo = fmaxf(a.y * vlo * vlo * 0.99999f - 1e-7f, 0.f);
        return false;
    }
    if (dmin >= 1.0f) { flo = fhi = 0.f; return true; }
    fhi = fminf(wyvill(dmin * dmin) * 1.00001f + 1e-7f, 1.0f);
    flo = fmaxf(wyvill(dmax * dmax) * 0.99999f - 1e-7f, 0.f);
    return false;
}


// prim_bound for ELONGATED boxes (used by the one-off slab prune only: in the per-region walks of the hierarchy its extra arithmetic
// cost more than the few primitives it drops - blob1k@512^3 cull 0.67 -> 0.88 ms with 0.7 % shorter programs).
// The distance to a skeleton is 1-Lipschitz, so over a ball (centre c, radius r) it lies in [d(c) - r, d(c) + r]. A cube-like box is
// covered by ONE ball (the lattice regions of every cull level); an elongated box - a rank's z-slab of 146 x 146 x 18 MPUs, whose single
// ball would reach across the whole model and keep every primitive, or the image of a region under a strong warp - is covered by up to
// 8 x 8 x 8 (at most 64; `fine`: 16 per axis, 1024) balls on a grid of cube-like tiles, and the bounds are the min / max over the tiles.
// `fine` (the one-off slab prune): up to 16 tiles per axis and 1024 in all, tiles of half the shortest side.
__device__ __noinline__ bool prim_bound_tiled(uint32_t hdr, const float4 q0, const float4* __restrict__ q, const Box3& b, float& flo, float& fhi, bool fine) {
    const uint32_t opc = PS_HDR_OPC(hdr);
    if (opc == OPC_NULLPRIM) { flo = fhi = 0.f; return true; }
    const float hx = 0.5f * (b.hi[0] - b.lo[0]), hy = 0.5f * (b.hi[1] - b.lo[1]), hz = 0.5f * (b.hi[2] - b.lo[2]);
    int nt[3] = {1, 1, 1};
    {
        const float hmin = fminf(hx, fminf(hy, hz)), hmax = fmaxf(hx, fmaxf(hy, hz));
        if (hmax > 1.5f * hmin && hmin > 0.f) {
            const float h[3] = {hx, hy, hz};
#pragma unroll
            for (int a = 0; a < 3; a++) nt[a] = (int)fminf(fine ? 16.0f : 8.0f, ceilf((fine ? 2.0f : 1.0f) * h[a] / hmin - 0.25f));
            while (nt[0] * nt[1] * nt[2] > (fine ? 1024 : 64)) {  // halve the finest axis until the cover has at most 64 tiles
                const int a = (nt[0] >= nt[1] && nt[0] >= nt[2]) ? 0 : (nt[1] >= nt[2] ? 1 : 2);
                nt[a] = (nt[a] + 1) >> 1;
            }
        }
    }
    if (!PS_HDR_XF(hdr) && prim_outside_box(opc, q0, q, b)) { flo = fhi = 0.f; return true; }
    const int nq = ps_base_quads(opc);
    const float4* m = q + (nq - 1);
    float scale = 1.0f;
    if (PS_HDR_XF(hdr)) {
        float fro = 0.f;
#pragma unroll
        for (int a = 0; a < 3; a++) fro += m[a].x * m[a].x + m[a].y * m[a].y + m[a].z * m[a].z;
        scale = sqrtf(fro) * 1.00001f;
    }
    const float tx = hx / (float)nt[0], ty = hy / (float)nt[1], tz = hz / (float)nt[2];
    const float r = sqrtf(tx * tx + ty * ty + tz * tz) * scale;
    const bool single = (nt[0] * nt[1] * nt[2] == 1);
    float dmin = 3.0e38f, dmax = 0.f;
    for (int ix = 0; ix < nt[0]; ix++)
        for (int iy = 0; iy < nt[1]; iy++)
            for (int iz = 0; iz < nt[2]; iz++) {
                float x[1], y[1], z[1];
                if (single) { x[0] = 0.5f * (b.lo[0] + b.hi[0]); y[0] = 0.5f * (b.lo[1] + b.hi[1]); z[0] = 0.5f * (b.lo[2] + b.hi[2]); }
                else { x[0] = b.lo[0] + (float)(2 * ix + 1) * tx; y[0] = b.lo[1] + (float)(2 * iy + 1) * ty; z[0] = b.lo[2] + (float)(2 * iz + 1) * tz; }
                if (PS_HDR_XF(hdr)) {
                    float ox[1], oy[1], oz[1];
                    xform_points<1>(m, x, y, z, ox, oy, oz);
                    x[0] = ox[0]; y[0] = oy[0]; z[0] = oz[0];
                }
                float d2[1];
                prim_d2<1>(opc, q0, q, x, y, z, d2);
                if (!(d2[0] == d2[0])) { flo = 0.f; fhi = 1.f; return false; }  // NaN: keep, unknown
                const float d = sqrtf(d2[0]);
                const float eps = 4e-6f * (fabsf(x[0]) + fabsf(y[0]) + fabsf(z[0]) + 1.0f) + 1e-5f * r + 1e-5f * d + (single ? 0.f : 1e-5f * (hx + hy + hz));
                dmin = fminf(dmin, fmaxf(d - r - eps, 0.f));
                dmax = fmaxf(dmax, d + r + eps);
            }
    if (opc == OPC_QUADRIC) {  // support radius R, law scale * (1 - d^2/R^2)^2 (scale > 0 is checked by the linearizer)
        const float4 a = q[0];
        if (dmin >= a.z) { flo = fhi = 0.f; return true; }
        const float vhi = fmaxf(1.0f - dmin * dmin * a.x, 0.f), vlo = fmaxf(1.0f - dmax * dmax * a.x, 0.f);
        fhi = a.y * vhi * vhi * 1.00001f + 1e-7f;
        flo = fmaxf(a.y * vlo * vlo * 0.99999f - 1e-7f, 0.f);
        return false;
    }
    if (dmin >= 1.0f) { flo = fhi = 0.f; return true; }
    fhi = fminf(wyvill(dmin * dmin) * 1.00001f + 1e-7f, 1.0f);
    flo = fmaxf(wyvill(dmax * dmax) * 0.99999f - 1e-7f, 0.f);
    return false;
}

__device__ __forceinline__ void bound_combine(uint32_t cmb, float& lo, float& hi, float vlo, float vhi) {
    const bool nonNegIn = lo >= 0.f && vlo >= 0.f && (cmb == CMB_ADD || cmb == CMB_RADD || cmb == CMB_MAX || cmb == CMB_MIN);
    switch (cmb) {
    case CMB_ADD: case CMB_RADD: lo += vlo; hi += vhi; break;
    case CMB_MAX: lo = fmaxf(lo, vlo); hi = fmaxf(hi, vhi); break;
    case CMB_MIN: lo = fminf(lo, vlo); hi = fminf(hi, vhi); break;
    case CMB_DIF: lo = fminf(lo, 1.0f - vhi); hi = fminf(hi, 1.0f - vlo); break;
    case CMB_RDIF: { const float a = fminf(vlo, 1.0f - hi), b = fminf(vhi, 1.0f - lo); lo = a; hi = b; } break;
    case CMB_RSDIF: {
        float a, b;
        imul(vlo, vhi, 1.0f - hi, 1.0f - lo, a, b);
        lo = a; hi = b;
    } break;
    default: {
        float a, b;
        imul(lo, hi, 1.0f - vhi, 1.0f - vlo, a, b);
        lo = a; hi = b;
    } break;
    }
    const float e = 2e-6f * (fabsf(lo) + fabsf(hi)) + 1e-7f;
    lo -= e; hi += e;
    // sums, maxima and minima of non-negative values are non-negative in floating point too: without this a sum with lower bound 0 went
    // to -1e-7 and the next all-zero union operand left a useless `max(top, 0)` instruction in the pruned program (3 % of them)
    if (nonNegIn) lo = fmaxf(lo, 0.f);
}

// Prunes a program against a region box. EMIT = false only measures. finish() returns the pruned length in quads including
// the END quad, or 0 if the region needs no evaluation.
// colorExact (PS_SimdPoly semantics): a blend that loses an all-zero operand leaves an OPC_CSCALE marker, see program.h.
// `limit` (EMIT only) = the length the measuring pass returned: the walk writes TENTATIVE code that a later zero-fold may
// rewind (intersection with 0, all-zero warp bracket, ...), and such code can run past the final length - into the slot of
// the next region's program in the pool, which another thread is writing. Nothing at or beyond `limit` survives, so it is
// simply not stored.
// `dflag` (EMIT, optional): one byte per emitted quad, 1 = first quad of an instruction. The warp-cooperative passes of
// cull.cu find the instruction headers of a parent program through it (one lane per quad position).
//
// The walk is a state machine fed one instruction at a time, so that a warp can compute the primitive bounds of 32 quad
// positions in parallel and let one lane step through the few instructions that matter (prune_walk below is the plain
// sequential driver; cull.cu has the cooperative one).
template <bool EMIT>
struct Pruner {
    float blo[PS_MAX_VALUE_STACK], bhi[PS_MAX_VALUE_STACK];
    uint32_t bstart[PS_MAX_VALUE_STACK];
    bool bzero[PS_MAX_VALUE_STACK];
    Box3 wbox[PS_MAX_POINT_STACK];
    uint32_t wstart[PS_MAX_POINT_STACK];
    int sp, wsp;
    uint32_t emit;
    Box3 box;
    uint32_t lastHdrPos, lastHdrVal;  // header quad of the last instruction emitted (for the colour-scale bits)
    float4 lastHdrQuad;
    uint32_t lastScale, scaleCount;   // position of the trailing OPC_CSCALE quad, if emit == lastScale + 1
    const float4* src;
    float4* dst;
    uint8_t* dflag;
    bool colorExact;
    uint32_t limit;

    __device__ void init(const float4* src_, const Box3& region, float4* dst_, uint8_t* dflag_, bool colorExact_, uint32_t limit_) {
        sp = -1; wsp = 0; emit = 0; box = region;
        lastHdrPos = 0xFFFFFFFFu; lastHdrVal = 0; lastHdrQuad = make_float4(0.f, 0.f, 0.f, 0.f);
        lastScale = 0xFFFFFFFFu; scaleCount = 0;
        src = src_; dst = dst_; dflag = dflag_; colorExact = colorExact_; limit = limit_;
    }
    __device__ __forceinline__ void put(float4 v, bool start) {
        if (EMIT && emit < limit) { dst[emit] = v; if (dflag) dflag[emit] = start ? 1 : 0; }
        emit++;
    }
    __device__ __forceinline__ float4 hdrquad(uint32_t h) { lastHdrPos = 0xFFFFFFFFu; return make_float4(__uint_as_float(h), 0.f, 0.f, 0.f); }
    __device__ __forceinline__ void rewind(uint32_t pos) {
        emit = pos;
        if (lastScale != 0xFFFFFFFFu && lastScale >= emit) lastScale = 0xFFFFFFFFu;
        if (lastHdrPos != 0xFFFFFFFFu && lastHdrPos >= emit) lastHdrPos = 0xFFFFFFFFu;
    }
    // colour-scale marker on the current top entry: first into the 4 header bits of the entry's last instruction (free for the
    // field-only passes), the overflow into an explicit OPC_CSCALE quad. cscale(a); cscale(b) == cscale(a + b).
    __device__ void cscale(uint32_t n) {
        if (!colorExact || n == 0) return;
        if (lastScale + 1 != emit && lastHdrPos != 0xFFFFFFFFu && sp >= 0 && lastHdrPos >= bstart[sp] && lastHdrPos < emit) {
            const uint32_t have = PS_HDR_CSC(lastHdrVal), room = PS_HDR_CSC_MAX - have, take = n < room ? n : room;
            if (take) {
                lastHdrVal += take << PS_HDR_CSC_SHIFT;
                lastHdrQuad.x = __uint_as_float(lastHdrVal);
                if (EMIT && lastHdrPos < limit) dst[lastHdrPos] = lastHdrQuad;
                n -= take;
            }
            if (n == 0) return;
        }
        if (lastScale + 1 == emit) scaleCount += n;
        else { lastScale = emit; scaleCount = n; emit++; }
        if (EMIT && lastScale < limit) {
            dst[lastScale] = make_float4(__uint_as_float(PS_HDR(OPC_CSCALE, 0, 0, 0)), __uint_as_float(scaleCount), 0.f, 0.f);
            if (dflag) dflag[lastScale] = 1;
        }
    }
    // copies instruction (pc, q0) with a possibly rewritten header
    __device__ void copy(uint32_t pc, const float4 q0, uint32_t newHdr, int nq) {
        const uint32_t o = PS_HDR_OPC(newHdr);
        // (a PRE count - program.h - scales the primitive's own colour before the combine: later steps owed to the ENTRY must not be
        //  added to it, they go to an explicit OPC_CSCALE quad)
        if ((o < OPC_FIRST_NONPRIM || o == OPC_COMBINE) && !PS_HDR_PRE(newHdr)) { lastHdrPos = emit; lastHdrVal = newHdr; lastHdrQuad = make_float4(__uint_as_float(newHdr), q0.y, q0.z, q0.w); }
        else lastHdrPos = 0xFFFFFFFFu;
        put(make_float4(__uint_as_float(newHdr), q0.y, q0.z, q0.w), true);
        for (int i = 1; i < nq; i++) put(src[pc + i], false);
    }
    // `n` colour-scale steps owed by all-zero blend operands (and the markers they carried) that the caller skipped since the
    // last instruction it fed: they apply to the current top entry if that is not itself all-zero
    __device__ __forceinline__ void skipped_scale(uint32_t n) {
        if (n && sp >= 0 && !bzero[sp]) cscale(n);
    }
    // true for an all-zero primitive the caller may skip instead of feeding it to step_prim: 0 added to / smooth-subtracted
    // from the top changes nothing but the colour weights; `scale` = the colour-scale steps to pass to skipped_scale
    // nonNeg (the program has no difference operators, so no value on the stack can be negative - context.cu sets it at upload): an
    // all-zero union operand, max(top, 0), is skippable too. Unions of one-primitive groups are most of a pruned blob program
    // (the peephole of step_op folds `PUSH p; COMBINE max` into `p (max)`), and at the next finer level most of them are zero.
    __device__ static __forceinline__ bool skippable(uint32_t hdr, bool pz, uint32_t& scale, bool nonNeg = false) {
        const uint32_t cmb = PS_HDR_CMB(hdr);
        if (!pz || (cmb != CMB_ADD && cmb != CMB_SDIF && !(nonNeg && cmb == CMB_MAX))) return false;
        scale = (cmb == CMB_ADD ? 1u : 0u) + (PS_HDR_PRE(hdr) ? 0u : PS_HDR_CSC(hdr));  // a PRE count scaled the dropped primitive's own colour
        return true;
    }

    // one primitive instruction; (pz, flo, fhi) = prim_bound over the CURRENT box
    __device__ void step_prim(uint32_t pc, const float4 q0, uint32_t hdr, bool pz, float flo, float fhi) {
        const uint32_t cmb = PS_HDR_CMB(hdr);
        const int nq = ps_instr_quads(hdr);
        if (cmb == CMB_PUSH) {
            sp++;
            bstart[sp] = emit; bzero[sp] = pz; blo[sp] = flo; bhi[sp] = fhi;
            if (!pz) copy(pc, q0, hdr, nq);
        } else if (bzero[sp]) {
            if (!pz && (cmb == CMB_ADD || cmb == CMB_MAX || cmb == CMB_RDIF || cmb == CMB_RSDIF)) {
                // 0 + f, max(0, f), min(f, 1 - 0) (f <= 1), f * (1 - 0): the primitive becomes the entry
                copy(pc, q0, PS_HDR_SET_CMB(hdr, CMB_PUSH) & ~PS_HDR_PRE_BIT, nq);  // alone on the stack, pre- and post-scale are the same
                bzero[sp] = false; blo[sp] = flo; bhi[sp] = fhi;
                if (cmb == CMB_ADD) cscale(1);
            }  // min(0,f) = 0, min(0,1-f) = 0 (f <= 1), 0*(1-f) = 0: stays zero
        } else if (pz) {
            if (cmb == CMB_MAX) {
                if (blo[sp] < 0.f) { put(hdrquad(PS_HDR(OPC_NULLPRIM, CMB_MAX, 0, 0)), true); blo[sp] = fmaxf(blo[sp], 0.f); }
            } else if (cmb == CMB_MIN) {
                if (blo[sp] >= 0.f) { rewind(bstart[sp]); bzero[sp] = true; blo[sp] = bhi[sp] = 0.f; }
                else { put(hdrquad(PS_HDR(OPC_NULLPRIM, CMB_MIN, 0, 0)), true); bhi[sp] = fminf(bhi[sp], 0.f); }
            } else if (cmb == CMB_DIF) {  // min(top, 1 - 0)
                if (bhi[sp] > 1.0f) { put(hdrquad(PS_HDR(OPC_CLAMP1, 0, 0, 0)), true); bhi[sp] = 1.0f; blo[sp] = fminf(blo[sp], 1.0f); }
            } else if (cmb == CMB_ADD) cscale(1);  // top + 0 (top * (1 - 0): nothing)
            else if (cmb == CMB_RSDIF || (cmb == CMB_RDIF && bhi[sp] <= 1.0f)) {  // 0 * (1 - top), min(0, 1 - top) with top <= 1
                rewind(bstart[sp]); bzero[sp] = true; blo[sp] = bhi[sp] = 0.f;
            } else if (cmb == CMB_RDIF) {  // min(0, 1 - top) can go negative
                put(hdrquad(PS_HDR(OPC_DIF0, 0, 0, 0)), true);
                const float a = fminf(0.f, 1.0f - bhi[sp]) - 1e-6f, b = fminf(0.f, 1.0f - blo[sp]);
                blo[sp] = a; bhi[sp] = b;
            }
            if (!bzero[sp] && !PS_HDR_PRE(hdr)) cscale(PS_HDR_CSC(hdr));  // markers a coarser level left on the dropped instruction stay with the entry
        } else {
            copy(pc, q0, hdr, nq);
            bound_combine(cmb, blo[sp], bhi[sp], flo, fhi);
        }
    }

    // T (cmb) 0 for a non-zero entry T = stack slot t (the slot of the all-zero operand is already popped)
    __device__ __forceinline__ void combine_with_zero(int t, uint32_t cmb, uint32_t hdr) {
        if (cmb == CMB_MAX) {
            if (blo[t] < 0.f) { put(hdrquad(PS_HDR(OPC_NULLPRIM, CMB_MAX, 0, 0)), true); blo[t] = 0.f; bhi[t] = fmaxf(bhi[t], 0.f); }
        } else if (cmb == CMB_MIN) {
            if (blo[t] >= 0.f) { rewind(bstart[t]); bzero[t] = true; blo[t] = bhi[t] = 0.f; }
            else { put(hdrquad(PS_HDR(OPC_NULLPRIM, CMB_MIN, 0, 0)), true); bhi[t] = fminf(bhi[t], 0.f); }
        } else if (cmb == CMB_DIF) {
            if (bhi[t] > 1.0f) { put(hdrquad(PS_HDR(OPC_CLAMP1, 0, 0, 0)), true); bhi[t] = 1.0f; blo[t] = fminf(blo[t], 1.0f); }
        } else if (cmb == CMB_ADD) cscale(1);
        if (!bzero[t]) cscale(PS_HDR_CSC(hdr));
    }
    // "PUSH <all-zero primitive>; OPC_COMBINE(CMB_MAX, no marker)" leaves emit, stack and markers untouched when the entry below is
    // all-zero (0 max 0) or cannot be negative (max(T, 0) = T): what step_prim + step_op would conclude, decided in a few instructions
    __device__ __forceinline__ bool zero_union_is_noop() const { return sp >= 0 && (bzero[sp] || blo[sp] >= 0.f); }
    // "PUSH <all-zero primitive>; OPC_COMBINE" with nothing fed in between, in one step: what step_prim + step_op would do,
    // without the stack round trip (the cooperative walks of cull.cu meet this pair once per culled blend group)
    __device__ __forceinline__ void step_zero_operand(uint32_t combineHdr) {
        const uint32_t cmbRaw = PS_HDR_CMB(combineHdr);
        const uint32_t cmb = (cmbRaw == CMB_RADD) ? (uint32_t)CMB_ADD : cmbRaw;
        if (sp >= 0 && !bzero[sp]) combine_with_zero(sp, cmb, combineHdr);
    }

    // one non-primitive instruction (not OPC_END); warps / operator matrices / OPC_POPPOINT change `box`
    __device__ void step_op(uint32_t pc, const float4 q0, uint32_t hdr) {
        const uint32_t opc = PS_HDR_OPC(hdr);
        const int nq = ps_instr_quads(hdr);
        switch (opc) {
        case OPC_COMBINE: {
            const uint32_t cmbRaw = PS_HDR_CMB(hdr);
            const uint32_t cmb = (cmbRaw == CMB_RADD) ? (uint32_t)CMB_ADD : cmbRaw;  // same zero-folding as a blend
            const int v = sp, t = sp - 1;
            sp--;
            if (bzero[t] && bzero[v]) {
                // nothing was emitted for either
            } else if (bzero[t]) {  // 0 (cmb) V; V's code starts at bstart[t]
                if (cmb == CMB_ADD) { bzero[t] = false; blo[t] = blo[v]; bhi[t] = bhi[v]; cscale(1); }
                else if (cmb == CMB_MAX) {
                    bzero[t] = false; blo[t] = blo[v]; bhi[t] = bhi[v];
                    if (blo[v] < 0.f) { put(hdrquad(PS_HDR(OPC_NULLPRIM, CMB_MAX, 0, 0)), true); blo[t] = 0.f; bhi[t] = fmaxf(bhi[v], 0.f); }
                } else if (cmb == CMB_MIN) {
                    if (blo[v] >= 0.f) rewind(bstart[t]);
                    else { put(hdrquad(PS_HDR(OPC_NULLPRIM, CMB_MIN, 0, 0)), true); bzero[t] = false; blo[t] = fminf(blo[v], 0.f); bhi[t] = 0.f; }
                } else if (cmb == CMB_DIF) {  // min(0, 1 - V)
                    if (bhi[v] <= 1.0f) rewind(bstart[t]);
                    else { put(hdrquad(PS_HDR(OPC_DIF0, 0, 0, 0)), true); bzero[t] = false; blo[t] = fminf(0.f, 1.0f - bhi[v]) - 1e-6f; bhi[t] = 0.f; }
                } else rewind(bstart[t]);  // 0 * (1 - V)
                if (!bzero[t]) cscale(PS_HDR_CSC(hdr));
            } else if (bzero[v]) {  // T (cmb) 0
                combine_with_zero(t, cmb, hdr);
            } else {
                // peephole: V shrank to ONE primitive (its blend partners were culled): fold it into T with the combine code in
                // its own header instead of PUSH + OPC_COMBINE - same operands in the same order, so the same bits, and the
                // interpreters save a stack round trip and a dispatch (most groups of a pruned list are single primitives)
                // If the primitive carries colour-scale steps of its own they become a PRE count (they scaled V's colour before the
                // combine); the OPC_COMBINE's own steps (after the combine) can ride along only if the primitive has none.
                const bool single = cmbRaw != CMB_RADD && lastHdrPos != 0xFFFFFFFFu && lastHdrPos == bstart[v] && lastScale + 1 != emit &&
                                    PS_HDR_OPC(lastHdrVal) < OPC_FIRST_NONPRIM && PS_HDR_CMB(lastHdrVal) == CMB_PUSH &&
                                    (PS_HDR_CSC(lastHdrVal) == 0u || PS_HDR_CSC(hdr) == 0u) &&
                                    emit == lastHdrPos + (uint32_t)ps_instr_quads(lastHdrVal);
                if (single) {
                    const bool pre = PS_HDR_CSC(lastHdrVal) != 0u;
                    lastHdrVal = PS_HDR_SET_CMB(lastHdrVal, cmb) | (pre ? PS_HDR_PRE_BIT : (PS_HDR_CSC(hdr) << PS_HDR_CSC_SHIFT));
                    lastHdrQuad.x = __uint_as_float(lastHdrVal);
                    if (EMIT && lastHdrPos < limit) dst[lastHdrPos] = lastHdrQuad;
                    if (pre) lastHdrPos = 0xFFFFFFFFu;  // its count field now means something else: later steps go to an OPC_CSCALE quad
                } else copy(pc, q0, hdr, nq);
                bound_combine(cmb, blo[t], bhi[t], blo[v], bhi[v]);
            }
        } break;
        case OPC_CSCALE: {
            if (!bzero[sp]) cscale(__float_as_uint(q0.y));
        } break;
        case OPC_POW: {
            if (!bzero[sp]) {
                copy(pc, q0, hdr, nq);
                const float e = q0.y;
                const float a = powf(fmaxf(blo[sp], 0.f), e), b = powf(fmaxf(bhi[sp], 0.f), e);
                blo[sp] = fmaxf(fminf(a, b) * 0.9999f - 1e-7f, 0.f); bhi[sp] = fmaxf(a, b) * 1.0001f + 1e-7f;  // a power of a clamped base is >= 0
                if (!(blo[sp] == blo[sp]) || !(bhi[sp] == bhi[sp])) { blo[sp] = -3.0e38f; bhi[sp] = 3.0e38f; }
            }
        } break;
        case OPC_RICCI_SIMD: {
            const int v = sp, t = sp - 1;
            sp--;
            if (bzero[t] && bzero[v]) {
                // fast_pow(0) = 0
            } else {
                if (bzero[t] || bzero[v]) put(hdrquad(PS_HDR(OPC_NULLPRIM, CMB_PUSH, 0, 0)), true);  // materialise the zero operand
                copy(pc, q0, hdr, nq);
                const float e = q0.y;
                const float b0 = (bzero[t] ? 0.f : blo[t]) + (bzero[v] ? 0.f : blo[v]);
                const float b1 = (bzero[t] ? 0.f : bhi[t]) + (bzero[v] ? 0.f : bhi[v]);
                const float a = ricci_simd(b0, e), b = ricci_simd(b1, e);
                bzero[t] = false;
                blo[t] = fminf(a, b) - 1e-5f * fabsf(a) - 1e-6f; bhi[t] = fmaxf(a, b) + 1e-5f * fabsf(b) + 1e-6f;
                if (!(e > 0.f) || !(blo[t] == blo[t]) || !(bhi[t] == bhi[t]) || b0 < 0.f) { blo[t] = -3.0e38f; bhi[t] = 3.0e38f; }
            }
        } break;
        case OPC_BOXMASK: {  // top, or 0 where the reference's deep-operator prune fires (program.h): an all-zero entry stays all-zero
            if (!bzero[sp]) { copy(pc, q0, hdr, nq); blo[sp] = fminf(blo[sp], 0.f); bhi[sp] = fmaxf(bhi[sp], 0.f); }
        } break;
        case OPC_CLAMP1: {
            if (!bzero[sp]) { copy(pc, q0, hdr, nq); bhi[sp] = fminf(bhi[sp], 1.0f); blo[sp] = fminf(blo[sp], 1.0f); }
        } break;
        case OPC_DIF0: {
            if (!bzero[sp]) {
                copy(pc, q0, hdr, nq);
                const float a = fminf(0.f, 1.0f - bhi[sp]) - 1e-6f, b = fminf(0.f, 1.0f - blo[sp]);
                blo[sp] = a; bhi[sp] = b;
            }
        } break;
        case OPC_POPPOINT: {
            wsp--;
            box = wbox[wsp];
            if (bzero[sp]) rewind(wstart[wsp]);
            else copy(pc, q0, hdr, nq);
            bstart[sp] = wstart[wsp];
        } break;
        default: {  // warps, operator affine
            wbox[wsp] = box;
            wstart[wsp] = emit;
            wsp++;
            copy(pc, q0, hdr, nq);
            warp_box(opc, q0, src + pc + 1, box);
        } break;
        }
    }

    __device__ uint32_t finish(int cullMode, float iso) {
        if (sp < 0 || bzero[0]) return 0;
        if (cullMode >= PSGPU_CULL_INTERVAL) {
            if (bhi[0] < iso - 1e-4f || blo[0] > iso + 1e-4f) return 0;
        }
        put(make_float4(__uint_as_float(PS_HDR(OPC_END, 0, 0, 0)), 0.f, 0.f, 0.f), true);
        return emit;
    }
};

// OPC_COMBINE header of a union that carries no colour-scale marker (or markers do not matter in this pass)
__device__ __forceinline__ bool ps_is_plain_union(uint32_t hdr, bool colorExact) {
#ifdef PSGPU_NO_ZEROPAIR  // A/B switch
    return false;
#else
    return PS_HDR_OPC(hdr) == OPC_COMBINE && PS_HDR_CMB(hdr) == CMB_MAX && (!colorExact || PS_HDR_CSC(hdr) == 0u);
#endif
}

// does this opcode change the region box of the walk (so that bounds of later primitives depend on it)?
__device__ __forceinline__ bool ps_changes_box(uint32_t opc) { return opc >= OPC_TWIST && opc <= OPC_POPPOINT; }

// plain sequential driver (one thread per region)
template <bool EMIT>
__device__ uint32_t prune_walk(const float4* __restrict__ src, const Box3& region, int cullMode, float iso, float4* __restrict__ dst,
                               bool colorExact = false, uint32_t limit = 0xFFFFFFFFu, uint8_t* dflag = nullptr, bool fineTiles = false) {
    Pruner<EMIT> P;
    P.init(src, region, dst, dflag, colorExact, limit);
    for (uint32_t pc = 0;;) {
        const float4 q0 = src[pc];
        const uint32_t hdr = __float_as_uint(q0.x);
        const uint32_t opc = PS_HDR_OPC(hdr);
        if (opc == OPC_END) break;
        if (opc < OPC_FIRST_NONPRIM) {
            float flo, fhi;
            const bool pz = fineTiles ? prim_bound_tiled(hdr, q0, src + pc + 1, P.box, flo, fhi, true) : prim_bound(hdr, q0, src + pc + 1, P.box, flo, fhi);
            P.step_prim(pc, q0, hdr, pz, flo, fhi);
        } else P.step_op(pc, q0, hdr);
        pc += ps_instr_quads(hdr);
    }
    return P.finish(cullMode, iso);
}

}  // namespace psgpu
#endif
