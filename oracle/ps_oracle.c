/* TEST INFRASTRUCTURE ONLY - see ps_oracle.h. Plain-C restatement of the reference's polygonization
 * path. SP/ = Parsip100/PS_SimdPoly/include/, PH/ = Parsip100/ParsipHaptics/include/,
 * FW/ = Parsip100/PS_FrameWork/include/ (all under /root/reference).
 *
 * Built with -O2 -msse2 -mfpmath=sse -ffp-contract=off so that every float expression rounds like the
 * reference's SSE code (no FMA contraction, no x87 excess precision); expression order is kept.
 */
#include "ps_oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "../parsip_b200/csrc/mc_tables.h"

static const int kTriTable[256 * 16] = PS_MC_TRITABLE_INIT;
static const int kCorner1[12] = PS_MC_CORNER1_INIT;
static const int kCorner2[12] = PS_MC_CORNER2_INIT;
static const int kEdgeAxis[12] = PS_MC_EDGEAXIS_INIT;

#define LANES 4 /* SIMD_USE_M128, SP/PS_SIMDVecN.h:17 */
#define GD PSO_GRID_DIM

/* _mm_max_ps(a,b) = a > b ? a : b ; _mm_min_ps(a,b) = a < b ? a : b (second operand on NaN) */
static inline float sse_max(float a, float b) { return a > b ? a : b; }
static inline float sse_min(float a, float b) { return a < b ? a : b; }

/* ---------------------------------------------------------------------------------------------
 * growable result mesh shared by both halves
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    float *pos, *nrm, *col;
    uint32_t* tri;
    int32_t* mpuIdx;
    uint32_t* mpuCounts;
    int64_t nv, nt, nm, capv, capt, capm;
} pso_mesh;

static void mesh_clear(pso_mesh* m) { m->nv = m->nt = m->nm = 0; }
static void mesh_free(pso_mesh* m) {
    free(m->pos); free(m->nrm); free(m->col); free(m->tri); free(m->mpuIdx); free(m->mpuCounts);
    memset(m, 0, sizeof(*m));
}
static void mesh_reserve(pso_mesh* m, int64_t dv, int64_t dt) {
    if (m->nv + dv > m->capv) {
        int64_t c = (m->nv + dv) * 2 + 1024;
        m->pos = (float*)realloc(m->pos, c * 3 * sizeof(float));
        m->nrm = (float*)realloc(m->nrm, c * 3 * sizeof(float));
        m->col = (float*)realloc(m->col, c * 3 * sizeof(float));
        m->capv = c;
    }
    if (m->nt + dt > m->capt) {
        int64_t c = (m->nt + dt) * 2 + 1024;
        m->tri = (uint32_t*)realloc(m->tri, c * 3 * sizeof(uint32_t));
        m->capt = c;
    }
    if (m->nm + 1 > m->capm) {
        int64_t c = (m->nm + 1) * 2 + 256;
        m->mpuIdx = (int32_t*)realloc(m->mpuIdx, c * 3 * sizeof(int32_t));
        m->mpuCounts = (uint32_t*)realloc(m->mpuCounts, c * 2 * sizeof(uint32_t));
        m->capm = c;
    }
}

/* one MPU's mesh; capacities are the worst case of a 7^3-cell block, not the reference's 512 */
#define MPU_MAXV (GD * GD * GD * 3)
#define MPU_MAXT (7 * 7 * 7 * 5)
typedef struct {
    float pos[MPU_MAXV * 3], nrm[MPU_MAXV * 3], col[MPU_MAXV * 3];
    uint16_t tri[MPU_MAXT * 3];
    int nv, nt;
    uint32_t fieldEvals;
    float lo[3];
} pso_mpu;

static void mesh_append(pso_mesh* m, const pso_mpu* u, const int* ijk) {
    mesh_reserve(m, u->nv, u->nt);
    memcpy(m->pos + m->nv * 3, u->pos, (size_t)u->nv * 3 * sizeof(float));
    memcpy(m->nrm + m->nv * 3, u->nrm, (size_t)u->nv * 3 * sizeof(float));
    memcpy(m->col + m->nv * 3, u->col, (size_t)u->nv * 3 * sizeof(float));
    for (int t = 0; t < u->nt * 3; t++) m->tri[m->nt * 3 + t] = (uint32_t)(u->tri[t] + m->nv);
    m->mpuIdx[m->nm * 3] = ijk[0]; m->mpuIdx[m->nm * 3 + 1] = ijk[1]; m->mpuIdx[m->nm * 3 + 2] = ijk[2];
    m->mpuCounts[m->nm * 2] = (uint32_t)u->nv; m->mpuCounts[m->nm * 2 + 1] = (uint32_t)u->nt;
    m->nv += u->nv; m->nt += u->nt; m->nm++;
}

/* MPU lattice: SP/PS_Polygonizer.cpp:328-352 == :388-412 (CountMPUNeeded) == PH/CPolyParsipOptimized.cpp:347-366 */
static void lattice_dims(float cellsize, const float* lo, const float* hi, int* nm) {
    const int cellsPerMPU = GD - 1;
    for (int a = 0; a < 3; a++) {
        float side = hi[a] - lo[a];
        int cells = (int)ceil(side / cellsize);
        nm[a] = cells / cellsPerMPU;
        if (cells % cellsPerMPU != 0) nm[a]++;
    }
}
uint32_t pso_count_mpus(float cellsize, const float* lo, const float* hi) {
    int nm[3];
    lattice_dims(cellsize, lo, hi, nm);
    return (uint32_t)(nm[0] * nm[1] * nm[2]);
}

/* =============================================================================================
 * SIMD-path semantics
 * =========================================================================================== */
struct pso_simd {
    int nP, nO, nM;
    int* ptype; float *ppos, *pdir, *pres, *pcol; int* pmat; float* pbox; /* n*6 lo,hi */
    int *otype, *oleft, *oright, *okind; float *ores, *obox;
    float *mbwd, *mfwd;
    float lo[3], hi[3];
    pso_mesh mesh;
};

static void* dup_mem(const void* p, size_t n) {
    void* q = malloc(n ? n : 1);
    if (p && n) memcpy(q, p, n); else if (n) memset(q, 0, n);
    return q;
}

pso_simd* pso_simd_create(const pso_soa_desc* d) {
    if (d->n_prims <= 0) return NULL;
    pso_simd* s = (pso_simd*)calloc(1, sizeof(pso_simd));
    s->nP = d->n_prims; s->nO = d->n_ops; s->nM = d->n_mats > 0 ? d->n_mats : 1;
    s->ptype = (int*)dup_mem(d->prim_type, sizeof(int) * s->nP);
    s->ppos = (float*)dup_mem(d->prim_pos, sizeof(float) * 3 * s->nP);
    s->pdir = (float*)dup_mem(d->prim_dir, sizeof(float) * 3 * s->nP);
    s->pres = (float*)dup_mem(d->prim_res, sizeof(float) * 3 * s->nP);
    s->pcol = (float*)dup_mem(d->prim_color, sizeof(float) * 3 * s->nP);
    s->pmat = (int*)dup_mem(d->prim_mat, sizeof(int) * s->nP);
    s->pbox = (float*)calloc((size_t)s->nP * 6, sizeof(float));
    s->otype = (int*)dup_mem(d->op_type, sizeof(int) * s->nO);
    s->oleft = (int*)dup_mem(d->op_left, sizeof(int) * s->nO);
    s->oright = (int*)dup_mem(d->op_right, sizeof(int) * s->nO);
    s->okind = (int*)dup_mem(d->op_kind, sizeof(int) * s->nO);
    s->ores = (float*)dup_mem(d->op_res, sizeof(float) * 4 * s->nO);
    s->obox = (float*)calloc((size_t)(s->nO ? s->nO : 1) * 6, sizeof(float));
    s->mbwd = (float*)calloc((size_t)s->nM * 12, sizeof(float));
    s->mfwd = (float*)calloc((size_t)s->nM * 16, sizeof(float));
    if (d->n_mats > 0 && d->mats_bwd) memcpy(s->mbwd, d->mats_bwd, sizeof(float) * 12 * d->n_mats);
    if (d->n_mats > 0 && d->mats_fwd) memcpy(s->mfwd, d->mats_fwd, sizeof(float) * 16 * d->n_mats);
    s->mbwd[0] = s->mbwd[5] = s->mbwd[10] = 1.0f;
    s->mfwd[0] = s->mfwd[5] = s->mfwd[10] = s->mfwd[15] = 1.0f;
    return s;
}

void pso_simd_destroy(pso_simd* s) {
    if (!s) return;
    free(s->ptype); free(s->ppos); free(s->pdir); free(s->pres); free(s->pcol); free(s->pmat); free(s->pbox);
    free(s->otype); free(s->oleft); free(s->oright); free(s->okind); free(s->ores); free(s->obox);
    free(s->mbwd); free(s->mfwd);
    mesh_free(&s->mesh);
    free(s);
}

void pso_simd_get_bbox(const pso_simd* s, float* lo, float* hi) { memcpy(lo, s->lo, 12); memcpy(hi, s->hi, 12); }
void pso_simd_set_bbox(pso_simd* s, const float* lo, const float* hi) { memcpy(s->lo, lo, 12); memcpy(s->hi, hi, 12); }
void pso_simd_get_prim_boxes(const pso_simd* s, float* out6) { memcpy(out6, s->pbox, sizeof(float) * 6 * s->nP); }
void pso_simd_get_op_boxes(const pso_simd* s, float* out6) { memcpy(out6, s->obox, sizeof(float) * 6 * s->nO); }
void pso_simd_pad_op_boxes(pso_simd* s, float pad) {
    for (int i = 0; i < s->nO; i++)
        for (int a = 0; a < 3; a++) { s->obox[i * 6 + a] -= pad; s->obox[i * 6 + 3 + a] += pad; }
}

/* FastSqrt, SP/PS_MathBase.h:233-247: inverse-sqrt bit trick + two Newton steps, times the number.
 * The reference reads the float through a `long*` (8 bytes on LP64, formally UB); the value it
 * produces in the unmodified build equals the 32-bit reading used here (= ISO_DIST 0.45420206f,
 * SP/PS_Polygonizer.h:59), checked bit-for-bit through the primitive boxes in tests/test_oracle_pinning.py. */
static float fast_sqrt(float number) {
    float x = number * 0.5F, y = number;
    int32_t i;
    memcpy(&i, &y, 4);
    i = 0x5f3759df - (i >> 1);
    memcpy(&y, &i, 4);
    y = y * (1.5F - (x * y * y));
    y = y * (1.5F - (x * y * y));
    return number * y;
}
/* ComputeWyvillIsoDistance, SP/PS_Polygonizer.h:419-424 */
static float wyvill_iso_distance(float fv) {
    if (fv == 0.0f) return 1.0f;
    float oneThird = 1.0f / 3.0f;
    return fast_sqrt(1.0f - powf(fv, oneThird));
}

/* mat4Transform of a point, SP/PS_MATRIX4.h:199-209: e[col][row] storage, translation in e[3][0..2] */
static void mat4_transform(const float* m, const float* v, float* out) {
    float r[3];
    for (int a = 0; a < 3; a++) r[a] = (m[0 * 4 + a] * v[0] + m[1 * 4 + a] * v[1] + m[2 * 4 + a] * v[2]) + m[3 * 4 + a];
    out[0] = r[0]; out[1] = r[1]; out[2] = r[2];
}

/* PrepareBBoxes, SP/PS_Polygonizer.cpp:55-309 */
int pso_simd_prepare_bboxes(pso_simd* s, float cellsize) {
    (void)cellsize;
    if (s->nP == 0) return PSGPU_PARAM_ERROR;
    float isoDist = wyvill_iso_distance(PSO_ISO_VALUE);
    isoDist += 5.0f * 0.01f; /* 5 * MIN_CELL_SIZE, :65 */
    float bl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, bh[3] = {FLT_MIN, FLT_MIN, FLT_MIN};
    for (int i = 0; i < s->nP; i++) {
        const float* p = &s->ppos[i * 3];
        const float* d = &s->pdir[i * 3];
        const float* r = &s->pres[i * 3];
        float* lo = &s->pbox[i * 6];
        float* hi = lo + 3;
        switch (s->ptype[i]) {
        case PSGPU_PRIM_POINT: /* :75-85 */
            for (int a = 0; a < 3; a++) { lo[a] = p[a] - isoDist; hi[a] = p[a] + isoDist; }
            break;
        case PSGPU_PRIM_LINE: /* :86-100 */
            for (int a = 0; a < 3; a++) {
                float expand = isoDist * 1.0f + (3.0f * isoDist) * (d[a] - p[a]);
                lo[a] = p[a] - expand; hi[a] = d[a] + expand;
            }
            break;
        case PSGPU_PRIM_RING: case PSGPU_PRIM_DISC: { /* :101-140 */
            float radius = r[0];
            radius += isoDist;
            for (int a = 0; a < 3; a++) {
                float dirComp = 1.0f - d[a];
                float expand = (radius + isoDist) * dirComp + isoDist * d[a];
                lo[a] = p[a] - expand; hi[a] = p[a] + expand;
            }
        } break;
        case PSGPU_PRIM_CYLINDER: { /* :141-159 */
            float radius = r[0], height = r[1];
            for (int a = 0; a < 3; a++) {
                float s1 = p[a] + height * d[a];
                float expand = (isoDist + radius) * 1.0f + (0.5f * isoDist) * d[a];
                lo[a] = p[a] - expand; hi[a] = s1 + expand;
            }
        } break;
        case PSGPU_PRIM_CUBE: { /* :160-176 */
            float side = r[0] + isoDist;
            for (int a = 0; a < 3; a++) { lo[a] = p[a] - side; hi[a] = p[a] + side; }
        } break;
        default: /* no case in the reference switch: the box keeps its previous content */
            break;
        }
        float curLo[3] = {lo[0], lo[1], lo[2]}, curHi[3] = {hi[0], hi[1], hi[2]};
        int im = s->pmat[i];
        if (im != 0) { /* :200-218, only the two corners are transformed */
            mat4_transform(&s->mfwd[im * 16], curLo, curLo);
            mat4_transform(&s->mfwd[im * 16], curHi, curHi);
            for (int a = 0; a < 3; a++) { lo[a] = curLo[a]; hi[a] = curHi[a]; }
        }
        if (i == 0) { memcpy(bl, curLo, 12); memcpy(bh, curHi, 12); }
        else for (int a = 0; a < 3; a++) { bl[a] = sse_min(bl[a], curLo[a]); bh[a] = sse_max(bh[a], curHi[a]); }
    }
    memcpy(s->lo, bl, 12); memcpy(s->hi, bh, 12);
    /* operator boxes = union of children, explicit-stack post-order, :238-307 */
    if (s->nO > 0) {
        int* stk = (int*)malloc(sizeof(int) * (size_t)(s->nO + 1) * 2);
        unsigned char* done = (unsigned char*)calloc((size_t)s->nO, 1);
        int top = -1;
        stk[++top] = 0;
        while (top >= 0) {
            int op = stk[top], lc = s->oleft[op], rc = s->oright[op];
            int lop = (s->okind[op] & 2) >> 1, rop = s->okind[op] & 1;
            int ready = !((lop && !done[lc]) || (rop && !done[rc]));
            if (ready) {
                top--;
                const float* lb = lop ? &s->obox[lc * 6] : &s->pbox[lc * 6];
                const float* rb = rop ? &s->obox[rc * 6] : &s->pbox[rc * 6];
                for (int a = 0; a < 3; a++) {
                    s->obox[op * 6 + a] = sse_min(lb[a], rb[a]);
                    s->obox[op * 6 + 3 + a] = sse_max(lb[3 + a], rb[3 + a]);
                }
                done[op] = 1;
            } else {
                if (lop && !done[lc]) stk[++top] = lc;
                if (rop && !done[rc]) stk[++top] = rc;
            }
        }
        free(stk); free(done);
    }
    return PSGPU_SUCCESS;
}

/* computePrimitiveField for one lane, SP/PS_Polygonizer.cpp:934-1179 + ComputeWyvillFieldValueSquare_
 * SP/PS_Polygonizer.h:397-407 */
static float simd_prim_field(const pso_simd* s, int ip, float pX, float pY, float pZ) {
    float ptX, ptY, ptZ, dist2 = 0.0f;
    int im = s->pmat[ip];
    if (im == 0) { ptX = pX; ptY = pY; ptZ = pZ; }
    else { /* :951-969 */
        const float* m = &s->mbwd[im * 12];
        ptX = m[0] * pX + m[1] * pY + m[2] * pZ + m[3];
        ptY = m[4] * pX + m[5] * pY + m[6] * pZ + m[7];
        ptZ = m[8] * pX + m[9] * pY + m[10] * pZ + m[11];
    }
    const float* pos = &s->ppos[ip * 3];
    const float* dir = &s->pdir[ip * 3];
    const float* res = &s->pres[ip * 3];
    switch (s->ptype[ip]) {
    case PSGPU_PRIM_POINT: { /* :975-983 */
        float dx = pos[0] - ptX, dy = pos[1] - ptY, dz = pos[2] - ptZ;
        dist2 = (dx * dx) + (dy * dy) + (dz * dz);
    } break;
    case PSGPU_PRIM_LINE: { /* :984-1011: projection is NOT clamped (infinite line) */
        float l0x = pos[0], l0y = pos[1], l0z = pos[2];
        float ldx = dir[0] - l0x, ldy = dir[1] - l0y, ldz = dir[2] - l0z;
        float ldd = ldx * ldx + ldy * ldy + ldz * ldz;
        float dx = ptX - l0x, dy = ptY - l0y, dz = ptZ - l0z;
        float delta = dx * ldx + dy * ldy + dz * ldz;
        delta = delta / ldd;
        dx = ptX - (l0x + delta * ldx);
        dy = ptY - (l0y + delta * ldy);
        dz = ptZ - (l0z + delta * ldz);
        dist2 = (dx * dx) + (dy * dy) + (dz * dz);
    } break;
    case PSGPU_PRIM_CYLINDER: { /* :1012-1039 */
        float px = ptX - pos[0], py = ptY - pos[1], pz = ptZ - pos[2];
        float radius = res[0], height = res[1];
        float y = px * dir[0] + py * dir[1] + pz * dir[2];
        float x = sse_max(0.0f, sqrtf(px * px + py * py + pz * pz - y * y) - radius);
        float mask = (y > 0.0f) ? 1.0f : 0.0f;
        y = mask * sse_max(0.0f, y - height) + (1.0f - mask) * y;
        dist2 = x * x + y * y;
    } break;
    case PSGPU_PRIM_CUBE: { /* :1059-1097 */
        float side = res[0], minusSide = -1.0f * res[0];
        float dif[3] = {ptX - pos[0], ptY - pos[1], ptZ - pos[2]};
        for (int a = 0; a < 3; a++) {
            float mm = (minusSide > dif[a]) ? 1.0f : 0.0f;
            float mp = (dif[a] > side) ? 1.0f : 0.0f;
            float delta = (dif[a] + side) * mm + (dif[a] - side) * mp;
            if (a == 0) dist2 = delta * delta; else dist2 += delta * delta;
        }
    } break;
    case PSGPU_PRIM_DISC: { /* :1099-1132; SimdRSqrt evaluated exactly here (Q7) */
        float dX = ptX - pos[0], dY = ptY - pos[1], dZ = ptZ - pos[2];
        float nX = dir[0], nY = dir[1], nZ = dir[2], radius = res[0];
        float dot = nX * dX + nY * dY + nZ * dZ;
        float dirX = dX - nX * dot, dirY = dY - nY * dot, dirZ = dZ - nZ * dot;
        dot = dirX * dirX + dirY * dirY + dirZ * dirZ;
        float rsdot = 1.0f / sqrtf(dot);
        dirX = dirX * rsdot; dirY = dirY * rsdot; dirZ = dirZ * rsdot;
        nX = radius * dirX - dX; nY = radius * dirY - dY; nZ = radius * dirZ - dZ;
        float mask = (radius * radius >= dot) ? 1.0f : 0.0f;
        dist2 = mask * (dX * dX + dY * dY + dZ * dZ - dot) + (1.0f - mask) * (nX * nX + nY * nY + nZ * nZ);
    } break;
    case PSGPU_PRIM_RING: { /* :1134-1173 */
        float dX = ptX - pos[0], dY = ptY - pos[1], dZ = ptZ - pos[2];
        float nX = dir[0], nY = dir[1], nZ = dir[2], radius = res[0];
        float dot = nX * dX + nY * dY + nZ * dZ;
        float dirX = dX - nX * dot, dirY = dY - nY * dot, dirZ = dZ - nZ * dot;
        dot = dirX * dirX + dirY * dirY + dirZ * dirZ;
        float mask = (dot == 0.0f) ? 1.0f : 0.0f;
        dot = 1.0f / sqrtf(dot);
        dirX = dirX * dot; dirY = dirY * dot; dirZ = dirZ * dot;
        nX = radius * dirX - dX; nY = radius * dirY - dY; nZ = radius * dirZ - dZ;
        dist2 = mask * (radius * radius + dX * dX + dY * dY + dZ * dZ) + (1.0f - mask) * (nX * nX + nY * nY + nZ * nZ);
    } break;
    default: /* bntPrimNull and anything without a case: dist2 stays 0 -> field 1 (reference quirk) */
        break;
    }
    float r1 = 1.0f - dist2;
    float f = (r1 * r1) * r1;
    return sse_max(0.0f, f);
}

/* scratch for one packet evaluation */
typedef struct {
    float* primF; /* nP * LANES */
    float* opF;   /* nO * LANES */
    unsigned char* opDone;
    int* stkId; int* stkDepth;
    float* opCol; /* nO * 3 (lane 0 only) */
} simd_scratch;

static void scratch_init(const pso_simd* s, simd_scratch* w) {
    w->primF = (float*)calloc((size_t)s->nP * LANES, sizeof(float));
    w->opF = (float*)calloc((size_t)(s->nO + 1) * LANES, sizeof(float));
    w->opDone = (unsigned char*)calloc((size_t)s->nO + 1, 1);
    w->stkId = (int*)malloc(sizeof(int) * (size_t)(s->nO + 2));
    w->stkDepth = (int*)malloc(sizeof(int) * (size_t)(s->nO + 2));
    w->opCol = (float*)calloc((size_t)(s->nO + 1) * 3, sizeof(float));
}
static void scratch_free(simd_scratch* w) {
    free(w->primF); free(w->opF); free(w->opDone); free(w->stkId); free(w->stkDepth); free(w->opCol);
}

/* FieldComputer::fieldValue for one packet of LANES points, SP/PS_Polygonizer.cpp:1184-1376 */
static void simd_field_packet(const pso_simd* s, simd_scratch* w, const float* pX, const float* pY,
                              const float* pZ, float* out) {
    for (int l = 0; l < LANES; l++) out[l] = 0.0f;
    if (s->nO > 0) {
        memset(w->opDone, 0, (size_t)s->nO);
        int top = -1;
        w->stkId[++top] = 0; w->stkDepth[top] = 0;
        while (top >= 0) {
            int op = w->stkId[top], depth = w->stkDepth[top];
            int lc = s->oleft[op], rc = s->oright[op];
            int lop = (s->okind[op] & 2) >> 1, rop = s->okind[op] & 1;
            if (depth > 3) { /* :1228-1252: prune iff, for every lane, outside the x AND y AND z slabs */
                const float* b = &s->obox[op * 6];
                int any = 0;
                for (int l = 0; l < LANES; l++) {
                    int in = (pX[l] >= b[0] && b[3] >= pX[l]) || (pY[l] >= b[1] && b[4] >= pY[l]) ||
                             (pZ[l] >= b[2] && b[5] >= pZ[l]);
                    any |= in;
                }
                if (!any) {
                    top--;
                    for (int l = 0; l < LANES; l++) { out[l] = 0.0f; w->opF[op * LANES + l] = 0.0f; }
                    w->opDone[op] = 1;
                    continue;
                }
            }
            int ready = !((lop && !w->opDone[lc]) || (rop && !w->opDone[rc]));
            if (ready) {
                top--;
                float L[LANES], R[LANES];
                for (int l = 0; l < LANES; l++) {
                    if (lop) L[l] = w->opF[lc * LANES + l];
                    else { L[l] = simd_prim_field(s, lc, pX[l], pY[l], pZ[l]); w->primF[lc * LANES + l] = L[l]; }
                }
                for (int l = 0; l < LANES; l++) {
                    if (rop) R[l] = w->opF[rc * LANES + l];
                    else { R[l] = simd_prim_field(s, rc, pX[l], pY[l], pZ[l]); w->primF[rc * LANES + l] = R[l]; }
                }
                for (int l = 0; l < LANES; l++) {
                    float o = out[l];
                    switch (s->otype[op]) {
                    case PSGPU_OP_BLEND: o = L[l] + R[l]; break;
                    case PSGPU_OP_RICCI: { /* :1289-1293 SimdPower = fast_pow, SP/PS_SIMDVecN.h:122-128
                                              base * rcp(e - e*base + base); rcp evaluated exactly (Q6) */
                        float base = L[l] + R[l], e = s->ores[op * 4 + 1];
                        float denom = e * base; denom = e - denom; denom = base + denom;
                        o = base * (1.0f / denom);
                    } break;
                    case PSGPU_OP_UNION: o = sse_max(L[l], R[l]); break;
                    case PSGPU_OP_INTERSECT: o = sse_min(L[l], R[l]); break;
                    case PSGPU_OP_DIF: o = sse_min(L[l], 1.0f - R[l]); break;
                    case PSGPU_OP_SMOOTHDIF: o = L[l] * (1.0f - R[l]); break;
                    case PSGPU_OP_WARP_BEND: case PSGPU_OP_WARP_TWIST: case PSGPU_OP_WARP_TAPER:
                    case PSGPU_OP_WARP_SHEAR: o = L[l]; break; /* :1318-1337 warps pass the left child through */
                    default: break; /* no case: outField keeps its previous value */
                    }
                    out[l] = o;
                    w->opF[op * LANES + l] = o;
                }
                w->opDone[op] = 1;
            } else {
                if (lop && !w->opDone[lc]) { w->stkId[++top] = lc; w->stkDepth[top] = depth + 1; }
                if (rop && !w->opDone[rc]) { w->stkId[++top] = rc; w->stkDepth[top] = depth + 1; }
            }
        }
    } else { /* :1356-1368 */
        for (int i = 0; i < s->nP; i++)
            for (int l = 0; l < LANES; l++) {
                float f = simd_prim_field(s, i, pX[l], pY[l], pZ[l]);
                out[l] = out[l] + f;
                w->primF[i * LANES + l] = f;
            }
    }
}

static float simd_field_bcast(const pso_simd* s, simd_scratch* w, float x, float y, float z) {
    float X[LANES], Y[LANES], Z[LANES], F[LANES];
    for (int l = 0; l < LANES; l++) { X[l] = x; Y[l] = y; Z[l] = z; }
    simd_field_packet(s, w, X, Y, Z, F);
    return F[0];
}

/* fieldValueAndColor, SP/PS_Polygonizer.cpp:1378-1551 (point broadcast to all lanes; lane 0 kept) */
static float simd_field_color(const pso_simd* s, simd_scratch* w, float x, float y, float z, float* col) {
    memset(w->primF, 0, sizeof(float) * (size_t)s->nP * LANES); /* reference: uninitialised stack */
    memset(w->opF, 0, sizeof(float) * (size_t)(s->nO + 1) * LANES);
    float f = simd_field_bcast(s, w, x, y, z);
    col[0] = col[1] = col[2] = 0.0f;
    if (s->nO > 0) {
        unsigned char* done = w->opDone;
        memset(done, 0, (size_t)s->nO);
        int top = -1;
        w->stkId[++top] = 0;
        while (top >= 0) {
            int op = w->stkId[top], lc = s->oleft[op], rc = s->oright[op];
            int lop = (s->okind[op] & 2) >> 1, rop = s->okind[op] & 1;
            int ready = !((lop && !done[lc]) || (rop && !done[rc]));
            if (ready) {
                top--;
                float cur = w->opF[op * LANES];
                float lf, rf, lcl[3], rcl[3];
                if (lop) { lf = w->opF[lc * LANES]; memcpy(lcl, &w->opCol[lc * 3], 12); }
                else { lf = w->primF[lc * LANES]; memcpy(lcl, &s->pcol[lc * 3], 12); }
                if (rop) { rf = w->opF[rc * LANES]; memcpy(rcl, &w->opCol[rc * 3], 12); }
                else { rf = w->primF[rc * LANES]; memcpy(rcl, &s->pcol[rc * 3], 12); }
                float o[3] = {col[0], col[1], col[2]};
                switch (s->otype[op]) {
                case PSGPU_OP_BLEND: case PSGPU_OP_RICCI: /* :1475-1487 */
                    lf = 2.0f * (0.5f + lf) - 1.0f;
                    rf = 2.0f * (0.5f + rf) - 1.0f;
                    for (int a = 0; a < 3; a++) o[a] = lf * lcl[a] + rf * rcl[a];
                    break;
                case PSGPU_OP_UNION: case PSGPU_OP_INTERSECT: /* :1489-1500 */
                    lf = (cur - lf == 0.0f) ? 1.0f : 0.0f;
                    rf = (cur - rf == 0.0f) ? 1.0f : 0.0f;
                    for (int a = 0; a < 3; a++) o[a] = lf * lcl[a] + rf * rcl[a];
                    break;
                case PSGPU_OP_DIF: case PSGPU_OP_SMOOTHDIF: /* :1502-1512 */
                    lf = (lf == cur) ? 1.0f : 0.0f;
                    rf = (1.0f - rf == cur) ? 1.0f : 0.0f;
                    for (int a = 0; a < 3; a++) o[a] = lf * lcl[a] + rf * rcl[a];
                    break;
                case PSGPU_OP_WARP_TWIST: case PSGPU_OP_WARP_TAPER: case PSGPU_OP_WARP_BEND:
                case PSGPU_OP_WARP_SHEAR:
                    for (int a = 0; a < 3; a++) o[a] = lcl[a];
                    break;
                default: break;
                }
                done[op] = 1;
                memcpy(col, o, 12);
                memcpy(&w->opCol[op * 3], o, 12);
            } else {
                if (lop && !done[lc]) w->stkId[++top] = lc;
                if (rop && !done[rc]) w->stkId[++top] = rc;
            }
        }
    } else {
        col[0] = s->pcol[0]; col[1] = s->pcol[1]; col[2] = s->pcol[2]; /* :1544-1546 */
    }
    return f;
}

/* FieldComputer::normal, SP/PS_Polygonizer.cpp:1598-1622; SimdNormalize's rsqrt evaluated exactly */
static void simd_normal(const pso_simd* s, simd_scratch* w, float x, float y, float z, float f, float delta, float* n) {
    float dinv = -1.0f / delta;
    float fx = simd_field_bcast(s, w, x + delta, y, z);
    float fy = simd_field_bcast(s, w, x, y + delta, z);
    float fz = simd_field_bcast(s, w, x, y, z + delta);
    float nx = (fx - f) * dinv, ny = (fy - f) * dinv, nz = (fz - f) * dinv;
    float inv = 1.0f / sqrtf(nx * nx + ny * ny + nz * nz);
    n[0] = nx * inv; n[1] = ny * inv; n[2] = nz * inv;
}

/* FieldComputer::computeRootNewtonRaphson, SP/PS_Polygonizer.cpp:1625-1681, with gradient() :1576-1596: every operand is a point
 * broadcast to all lanes, lane 0 is returned. The field at the start point is the caller's fp1 / fp2, not re-evaluated. */
void pso_simd_newton(const pso_simd* s, int64_t n, const float* p1, const float* p2, const float* fp1, const float* fp2,
                     float target, int iterations, float* root, float* root_field, int32_t* evals) {
    simd_scratch w;
    scratch_init(s, &w);
    const float delta = 0.001f; /* FIELDVALUE_EPSILON, SP/PS_Polygonizer.h:60 */
    const float deltaInv = 1.0f / delta;
    for (int64_t k = 0; k < n; k++) {
        if (iterations <= 0) { evals[k] = -1; continue; }
        float x, y, z, field;
        if (fabsf(fp1[k] - target) < fabsf(fp2[k] - target)) { x = p1[k * 3]; y = p1[k * 3 + 1]; z = p1[k * 3 + 2]; field = fp1[k]; }
        else { x = p2[k * 3]; y = p2[k * 3 + 1]; z = p2[k * 3 + 2]; field = fp2[k]; }
        int i = 0;
        for (i = 0; i < iterations; i++) {
            float dX = simd_field_bcast(s, &w, x + delta, y, z);
            float dY = simd_field_bcast(s, &w, x, y + delta, z);
            float dZ = simd_field_bcast(s, &w, x, y, z + delta);
            float gx = (dX - field) * deltaInv, gy = (dY - field) * deltaInv, gz = (dZ - field) * deltaInv;
            float den = gx * gx; den = den + gy * gy; den = den + gz * gz;
            float g = (target - field) / den;
            x = x + gx * g; y = y + gy * g; z = z + gz * g;
            field = simd_field_bcast(s, &w, x, y, z);
            if (fabsf(field - target) < delta) break;
        }
        root[k * 3] = x; root[k * 3 + 1] = y; root[k * 3 + 2] = z;
        root_field[k] = field;
        evals[k] = (i + 1) * 4;
    }
    scratch_free(&w);
}

void pso_simd_field(const pso_simd* s, int64_t n, const float* xyz, float* out) {
    simd_scratch w;
    scratch_init(s, &w);
    for (int64_t b = 0; b < n; b += LANES) {
        float X[LANES], Y[LANES], Z[LANES], F[LANES];
        for (int l = 0; l < LANES; l++) {
            int64_t i = b + l < n ? b + l : n - 1;
            X[l] = xyz[i * 3]; Y[l] = xyz[i * 3 + 1]; Z[l] = xyz[i * 3 + 2];
        }
        simd_field_packet(s, &w, X, Y, Z, F);
        for (int l = 0; l < LANES && b + l < n; l++) out[b + l] = F[l];
    }
    scratch_free(&w);
}

void pso_simd_field_color_normal(const pso_simd* s, int64_t n, const float* xyz, float* f, float* col, float* nrm) {
    simd_scratch w;
    scratch_init(s, &w);
    for (int64_t i = 0; i < n; i++) {
        f[i] = simd_field_color(s, &w, xyz[i * 3], xyz[i * 3 + 1], xyz[i * 3 + 2], &col[i * 3]);
        simd_normal(s, &w, xyz[i * 3], xyz[i * 3 + 1], xyz[i * 3 + 2], f[i], PSO_NORMAL_DELTA, &nrm[i * 3]);
    }
    scratch_free(&w);
}

/* phase 1 of process_cells_simd: the 8x8x8 field cache, SP/PS_Polygonizer.cpp:562-600.
 * Returns 0 if the MPU is discarded (:602-609), else 1. */
static int simd_fill_cache(const pso_simd* s, simd_scratch* w, const float* lo, float cellsize, float* fv, uint32_t* evals) {
    float inside[LANES] = {0, 0, 0, 0}, outside[LANES] = {0, 0, 0, 0};
    for (int i = 0; i < GD; i++)
        for (int j = 0; j < GD; j++)
            for (int k = 0; k < GD / LANES; k++) {
                float X[LANES], Y[LANES], Z[LANES], F[LANES];
                float cx = lo[0] + i * cellsize, cy = lo[1] + j * cellsize;
                float kf = (float)k * LANES;
                for (int l = 0; l < LANES; l++) { X[l] = cx; Y[l] = cy; Z[l] = lo[2] + (kf + (float)l) * cellsize; }
                simd_field_packet(s, w, X, Y, Z, F);
                for (int l = 0; l < LANES; l++) {
                    fv[i * GD + j * GD * GD + k * LANES + l] = F[l];
                    float in = (F[l] >= PSO_ISO_VALUE) ? 1.0f : 0.0f;
                    inside[l] = inside[l] + in;
                    outside[l] = outside[l] + 1.0f - in;
                }
                (*evals)++;
            }
    int anyIn = 0, anyOut = 0;
    for (int l = 0; l < LANES; l++) { anyIn |= inside[l] != 0.0f; anyOut |= outside[l] != 0.0f; }
    return anyIn && anyOut;
}

static int simd_cell_case(const float* fv, int i, int j, int k) { /* :647-689 */
    int ip0 = i * GD, ip1 = ip0 + GD, jp0 = j * GD * GD, jp1 = jp0 + GD * GD;
    float c[8] = {fv[ip0 + jp0 + k], fv[ip0 + jp0 + k + 1], fv[ip0 + jp1 + k], fv[ip0 + jp1 + k + 1],
                  fv[ip1 + jp0 + k], fv[ip1 + jp0 + k + 1], fv[ip1 + jp1 + k], fv[ip1 + jp1 + k + 1]};
    int idx = 0;
    for (int b = 0; b < 8; b++) if (c[b] >= PSO_ISO_VALUE) idx |= 1 << b;
    return idx;
}

/* CMPUProcessor::process_cells_simd, SP/PS_Polygonizer.cpp:475-829 */
static void simd_process_mpu(const pso_simd* s, simd_scratch* w, float cellsize, pso_mpu* mpu) {
    float fv[GD * GD * GD];
    mpu->nv = mpu->nt = 0; mpu->fieldEvals = 0;
    { /* phase 0, :483-540: the MPU's 8 corners, reject iff all exactly 0 */
        float side = (GD - 1) * cellsize;
        int any = 0;
        for (int half = 0; half < 2; half++) {
            float X[LANES], Y[LANES], Z[LANES], F[LANES];
            for (int l = 0; l < LANES; l++) {
                int c = half * 4 + l;
                X[l] = ((float)((0xAA >> c) & 1) * side) + mpu->lo[0];
                Y[l] = ((float)((0xCC >> c) & 1) * side) + mpu->lo[1];
                Z[l] = ((float)((0xF0 >> c) & 1) * side) + mpu->lo[2];
            }
            simd_field_packet(s, w, X, Y, Z, F);
            for (int l = 0; l < LANES; l++) any |= (F[l] > 0.0f);
        }
        if (!any) return;
    }
    unsigned char hasEdge[GD * GD * GD * 3];
    uint16_t edgeVid[GD * GD * GD * 3];
    memset(hasEdge, 0, sizeof(hasEdge));
    if (!simd_fill_cache(s, w, mpu->lo, cellsize, fv, &mpu->fieldEvals)) return;
    float rootRes[LANES];
    for (int l = 0; l < LANES; l++) rootRes[l] = (float)l * (1.0f / 3.0f); /* :629-634 */
    for (int i = 0; i < GD - 1; i++)
        for (int j = 0; j < GD - 1; j++)
            for (int k = 0; k < GD - 1; k++) {
                int cfg = simd_cell_case(fv, i, j, k);
                if (cfg == 0 || cfg == 255) continue;
                const int* row = &kTriTable[cfg * 16];
                int vids[16], nEdges = 0;
                for (int ic = 0; ic < 16; ic++) {
                    int cand = row[ic];
                    if (cand == -1) break;
                    int es = kCorner1[cand], ax = kEdgeAxis[cand];
                    int sx = i + ((es & 4) >> 2), sy = j + ((es & 2) >> 1), sz = k + (es & 1);
                    int h = ((sz & 7) << 6) | ((sy & 7) << 3) | (sx & 7); /* CELLID_FROM_IDX, SP/PS_Polygonizer.h:53 */
                    if (hasEdge[h * 3 + ax]) vids[ic] = edgeVid[h * 3 + ax];
                    else { /* :719-808 */
                        float e1[3] = {mpu->lo[0] + cellsize * sx, mpu->lo[1] + cellsize * sy, mpu->lo[2] + cellsize * sz};
                        float e2[3] = {e1[0], e1[1], e1[2]};
                        e2[ax] = e2[ax] + cellsize;
                        float dX = e2[0] - e1[0], dY = e2[1] - e1[1], dZ = e2[2] - e1[2];
                        float X[LANES], Y[LANES], Z[LANES], F[LANES];
                        for (int l = 0; l < LANES; l++) {
                            X[l] = e1[0] + dX * rootRes[l]; Y[l] = e1[1] + dY * rootRes[l]; Z[l] = e1[2] + dZ * rootRes[l];
                        }
                        simd_field_packet(s, w, X, Y, Z, F);
                        int state = F[0] >= PSO_ISO_VALUE, interval = 0;
                        for (int l = 1; l < LANES; l++) { interval = l; if ((F[l] >= PSO_ISO_VALUE) != state) break; }
                        float a1[3] = {X[interval - 1], Y[interval - 1], Z[interval - 1]};
                        float a2[3] = {X[interval], Y[interval], Z[interval]};
                        float scale = (PSO_ISO_VALUE - F[interval - 1]) / (F[interval] - F[interval - 1]);
                        float p[3];
                        for (int a = 0; a < 3; a++) p[a] = a1[a] + scale * (a2[a] - a1[a]);
                        int v = mpu->nv;
                        float f = simd_field_color(s, w, p[0], p[1], p[2], &mpu->col[v * 3]);
                        simd_normal(s, w, p[0], p[1], p[2], f, PSO_NORMAL_DELTA, &mpu->nrm[v * 3]);
                        memcpy(&mpu->pos[v * 3], p, 12);
                        mpu->nv++;
                        vids[ic] = v;
                        hasEdge[h * 3 + ax] = 1; edgeVid[h * 3 + ax] = (uint16_t)v;
                    }
                    nEdges++;
                }
                for (int t = 0; t < nEdges / 3; t++) { /* :816-825 */
                    mpu->tri[mpu->nt * 3 + 0] = (uint16_t)vids[t * 3 + 0];
                    mpu->tri[mpu->nt * 3 + 1] = (uint16_t)vids[t * 3 + 1];
                    mpu->tri[mpu->nt * 3 + 2] = (uint16_t)vids[t * 3 + 2];
                    mpu->nt++;
                }
            }
}

void pso_simd_mpu_cache(const pso_simd* s, float cellsize, const int32_t* ijk, float* fv512, uint8_t* case343) {
    simd_scratch w;
    scratch_init(s, &w);
    float side = cellsize * (float)(GD - 1);
    float lo[3] = {s->lo[0] + (float)ijk[0] * side, s->lo[1] + (float)ijk[1] * side, s->lo[2] + (float)ijk[2] * side};
    uint32_t ev = 0;
    simd_fill_cache(s, &w, lo, cellsize, fv512, &ev);
    for (int i = 0; i < GD - 1; i++) for (int j = 0; j < GD - 1; j++) for (int k = 0; k < GD - 1; k++)
        case343[i * 49 + j * 7 + k] = (uint8_t)simd_cell_case(fv512, i, j, k);
    scratch_free(&w);
}

/* ---- parallel driver shared by both halves ---- */
typedef void (*mpu_fn)(const void* scene, void* scratch, float cellsize, pso_mpu* mpu);
typedef struct {
    const void* scene; mpu_fn fn; float cellsize; const float* lo;
    int B[3], E[3]; int64_t total; volatile int64_t* next; int keep;
    pthread_mutex_t* lock; pso_mesh* mesh; int64_t* stats; int isSimd;
} drv_ctx;

static void simd_mpu_tramp(const void* sc, void* scratch, float cellsize, pso_mpu* mpu) {
    simd_process_mpu((const pso_simd*)sc, (simd_scratch*)scratch, cellsize, mpu);
}

typedef struct { int64_t key; pso_mpu* mpu; } kept_mpu;

static void* drv_worker(void* arg);

static int run_driver(const void* scene, int isSimd, mpu_fn fn, float cellsize, const float* lo, const float* hi,
                      int nthreads, const int32_t* b0, const int32_t* e0, int keep, pso_mesh* mesh, int64_t* stats);

/* Polygonize, SP/PS_Polygonizer.cpp:315-385: x outer, z inner, origin = lo + idx*side */
int pso_simd_polygonize(pso_simd* s, float cellsize, int nthreads, const int32_t* b0, const int32_t* e0,
                        int keep_mesh, int64_t* stats) {
    if (s->nP == 0) return PSGPU_PARAM_ERROR;
    return run_driver(s, 1, simd_mpu_tramp, cellsize, s->lo, s->hi, nthreads, b0, e0, keep_mesh, &s->mesh, stats);
}
void pso_simd_result_counts(const pso_simd* s, int64_t* o) { o[0] = s->mesh.nv; o[1] = s->mesh.nt; o[2] = s->mesh.nm; }
static void result_copy(const pso_mesh* m, float* pos, float* nrm, float* col, uint32_t* tri, int32_t* mi, uint32_t* mc) {
    if (pos) memcpy(pos, m->pos, (size_t)m->nv * 12);
    if (nrm) memcpy(nrm, m->nrm, (size_t)m->nv * 12);
    if (col) memcpy(col, m->col, (size_t)m->nv * 12);
    if (tri) memcpy(tri, m->tri, (size_t)m->nt * 12);
    if (mi) memcpy(mi, m->mpuIdx, (size_t)m->nm * 12);
    if (mc) memcpy(mc, m->mpuCounts, (size_t)m->nm * 8);
}
void pso_simd_result_copy(const pso_simd* s, float* pos, float* nrm, float* col, uint32_t* tri, int32_t* mi, uint32_t* mc) {
    result_copy(&s->mesh, pos, nrm, col, tri, mi, mc);
}

/* =============================================================================================
 * scalar-twin semantics
 * =========================================================================================== */
struct pso_scalar {
    int n, root;
    int *type, *first, *nkids, *child;
    float *pos, *dir, *res, *col, *xf;
    unsigned char* hasxf;
    int nChild;
    pso_mesh mesh;
    float lo[3], hi[3];
    float *octLo, *octHi; /* per node, optional: the scene API's primitive octrees for the MPU filter */
};

pso_scalar* pso_scalar_create(const psgpu_tree* t) {
    if (t->n_nodes <= 0) return NULL;
    pso_scalar* s = (pso_scalar*)calloc(1, sizeof(pso_scalar));
    int n = t->n_nodes;
    s->n = n; s->root = t->root;
    int nc = 0;
    for (int i = 0; i < n; i++) if (t->first_child[i] + t->n_children[i] > nc) nc = t->first_child[i] + t->n_children[i];
    s->nChild = nc;
    s->type = (int*)dup_mem(t->type, sizeof(int) * n);
    s->first = (int*)dup_mem(t->first_child, sizeof(int) * n);
    s->nkids = (int*)dup_mem(t->n_children, sizeof(int) * n);
    s->child = (int*)dup_mem(t->child, sizeof(int) * (nc ? nc : 1));
    s->pos = (float*)dup_mem(t->pos, sizeof(float) * 3 * n);
    s->dir = (float*)dup_mem(t->dir, sizeof(float) * 3 * n);
    s->res = (float*)dup_mem(t->res, sizeof(float) * 4 * n);
    s->col = (float*)dup_mem(t->color, sizeof(float) * 3 * n);
    s->xf = (float*)dup_mem(t->xform, sizeof(float) * 12 * n);
    s->hasxf = (unsigned char*)dup_mem(t->has_xform, (size_t)n);
    if (!t->xform || !t->has_xform) memset(s->hasxf, 0, (size_t)n);
    return s;
}
void pso_scalar_destroy(pso_scalar* s) {
    if (!s) return;
    free(s->type); free(s->first); free(s->nkids); free(s->child); free(s->pos); free(s->dir); free(s->res);
    free(s->col); free(s->xf); free(s->hasxf); free(s->octLo); free(s->octHi); mesh_free(&s->mesh); free(s);
}

/* The scalar polygonizer skips an MPU unless its box overlaps some primitive's octree, PH/CPolyParsipOptimized.cpp:160-183
 * (strict overlap test :117-127). The octrees are the scene API's heuristic bounds (CSkeleton::bound(), e.g. BT/CSkeletonLine.h:80-85
 * shrinks the box of a line whose end is below its start), so they are an INPUT here: one box per node of the flat
 * tree, read for primitive nodes only. Without boxes every MPU is processed. */
void pso_scalar_set_prim_boxes(pso_scalar* s, const float* lo3n, const float* hi3n) {
    free(s->octLo); free(s->octHi); s->octLo = s->octHi = NULL;
    if (!lo3n || !hi3n) return;
    s->octLo = (float*)malloc(sizeof(float) * 3 * (size_t)s->n); s->octHi = (float*)malloc(sizeof(float) * 3 * (size_t)s->n);
    memcpy(s->octLo, lo3n, sizeof(float) * 3 * (size_t)s->n); memcpy(s->octHi, hi3n, sizeof(float) * 3 * (size_t)s->n);
}

/* ComputeWyvillFieldValueSquare, SP/PS_Polygonizer.h:410-416 == BT/CFieldFunction.h:104-114 */
static float wyvill_sq(float dd) {
    if (dd >= 1.0f) return 0.0f;
    float t = 1.0f - dd;
    return t * t * t;
}

/* warps, PH/CompactBlobTree.cpp:1315-1462 */
static void warp_bend(const float* pin, float k, float y0, float left, float right, float* pout) {
    float kDiv = 1.0f / k, yh = 0.0f;
    if (pin[1] <= left) yh = left;
    else if (pin[1] > left && pin[1] < right) yh = pin[1];
    else if (pin[1] >= right) yh = right;
    float theta = k * (yh - y0);
    float ct = cosf(theta), st = sinf(theta);
    int inside = (pin[1] >= left) && (pin[1] <= right); /* CInterval::isInside, FW/PS_Interval.h:50 */
    float oy = 0.0f, oz = 0.0f;
    if (inside) oy = -st * (pin[2] - kDiv) + y0;
    else if (pin[1] < left) oy = -st * (pin[2] - kDiv) + y0 + ct * (pin[1] - left);
    else if (pin[1] > right) oy = -st * (pin[2] - kDiv) + y0 + ct * (pin[1] - right);
    if (inside) oz = ct * (pin[2] - kDiv) + kDiv;
    else if (pin[1] < left) oz = ct * (pin[2] - kDiv) + kDiv + st * (pin[1] - left);
    else if (pin[1] > right) oz = ct * (pin[2] - kDiv) + kDiv + st * (pin[1] - right);
    pout[0] = pin[0]; pout[1] = oy; pout[2] = oz;
}
static void warp_twist(const float* pin, float factor, int axis, float* pout) {
    float x = pin[0], y = pin[1], z = pin[2], theta;
    switch (axis) {
    case 0: theta = x * factor; pout[0] = x; pout[1] = y * cosf(theta) - z * sinf(theta); pout[2] = y * sinf(theta) + z * cosf(theta); break;
    case 1: theta = y * factor; pout[0] = x * cosf(theta) - z * sinf(theta); pout[1] = y; pout[2] = x * sinf(theta) + z * cosf(theta); break;
    default: theta = z * factor; pout[0] = x * cosf(theta) - y * sinf(theta); pout[1] = x * sinf(theta) + y * cosf(theta); pout[2] = z; break;
    }
}
static void warp_taper(const float* p, float factor, int along, int taper, float* o) {
    o[0] = p[0]; o[1] = p[1]; o[2] = p[2];
    switch (along) {
    case 0: /* taperAlongX :1387-1403 */
        if (taper == 2) o[2] = p[2] * (1 + p[0] * factor); else o[1] = p[1] * (1 + p[0] * factor);
        break;
    case 1: /* taperAlongY :1405-1421 */
        if (taper == 2) o[2] = p[2] * (1 + p[1] * factor); else o[0] = p[0] * (1 + p[1] * factor);
        break;
    default: /* taperAlongZ :1423-1439: case zAxis tapers y (the reference's axis mix-up) */
        if (taper == 2) o[1] = p[1] * (1 + p[2] * factor); else o[0] = p[0] * (1 + p[2] * factor);
        break;
    }
}
static void warp_shear(const float* p, float factor, int along, int dep, float* o) {
    o[0] = p[0]; o[1] = p[1]; o[2] = p[2];
    switch (along) { /* :1464-1536 */
    case 1: o[1] = p[1] + factor * (dep == 2 ? p[2] : p[0]); break;
    case 2: o[2] = p[2] + factor * (dep == 1 ? p[1] : p[0]); break;
    default: o[0] = p[0] + factor * (dep == 2 ? p[2] : p[1]); break;
    }
}

static float scalar_node(const pso_scalar* s, int id, const float* p, float w, float* store);

/* fieldvaluePrim, PH/CompactBlobTree.cpp:893-1092 */
static float scalar_node(const pso_scalar* s, int id, const float* p, float w, float* store);

static float scalar_prim(const pso_scalar* s, int id, const float* p) {
    float pn[3] = {p[0], p[1], p[2]};
    if (s->hasxf[id]) {
        const float* m = &s->xf[id * 12]; /* row.dot(vec4(p,1)) :899-905 */
        pn[0] = m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3] * 1.0f;
        pn[1] = m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7] * 1.0f;
        pn[2] = m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11] * 1.0f;
    }
    const float* c = &s->pos[id * 3];
    const float* d = &s->dir[id * 3];
    const float* r = &s->res[id * 4];
    switch (s->type[id]) {
    case PSGPU_PRIM_POINT: { /* :910-914 */
        float dx = pn[0] - c[0], dy = pn[1] - c[1], dz = pn[2] - c[2];
        return wyvill_sq(dx * dx + dy * dy + dz * dz);
    }
    case PSGPU_PRIM_CYLINDER: { /* :915-930 */
        float px = pn[0] - c[0], py = pn[1] - c[1], pz = pn[2] - c[2];
        float y = px * d[0] + py * d[1] + pz * d[2];
        float x = fmaxf(0.0f, sqrtf((px * px + py * py + pz * pz) - y * y) - r[0]);
        if (y > 0.0f) y = fmaxf(0.0f, y - r[1]);
        return wyvill_sq(x * x + y * y);
    }
    case PSGPU_PRIM_CUBE: { /* :944-994 */
        float side = r[0], dist2 = 0.0f;
        for (int a = 0; a < 3; a++) {
            float proj = pn[a] - c[a];
            if (proj < -1.0f * side) { float dl = proj + side; dist2 += dl * dl; }
            else if (proj > side) { float dl = proj - side; dist2 += dl * dl; }
        }
        return wyvill_sq(dist2);
    }
    case PSGPU_PRIM_DISC: { /* :996-1017 */
        float q[3] = {pn[0] - c[0], pn[1] - c[1], pn[2] - c[2]};
        float nd = d[0] * q[0] + d[1] * q[1] + d[2] * q[2];
        float dir[3] = {q[0] - nd * d[0], q[1] - nd * d[1], q[2] - nd * d[2]};
        float dl2 = dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2];
        float dd;
        if (sqrtf(dl2) <= r[0]) dd = fabsf((q[0] * q[0] + q[1] * q[1] + q[2] * q[2]) - dl2);
        else {
            float inv = 1.0f / sqrtf(dl2);
            float x[3] = {c[0] + r[0] * (dir[0] * inv), c[1] + r[0] * (dir[1] * inv), c[2] + r[0] * (dir[2] * inv)};
            dd = (x[0] - pn[0]) * (x[0] - pn[0]) + (x[1] - pn[1]) * (x[1] - pn[1]) + (x[2] - pn[2]) * (x[2] - pn[2]);
        }
        return wyvill_sq(dd);
    }
    case PSGPU_PRIM_RING: { /* :1019-1040 */
        float q[3] = {pn[0] - c[0], pn[1] - c[1], pn[2] - c[2]};
        float nd = d[0] * q[0] + d[1] * q[1] + d[2] * q[2];
        float dir[3] = {q[0] - nd * d[0], q[1] - nd * d[1], q[2] - nd * d[2]};
        float dl2 = dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2];
        float dd;
        if (dir[0] == 0.0f && dir[1] == 0.0f && dir[2] == 0.0f) dd = r[0] * r[0] + (q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
        else {
            float inv = 1.0f / sqrtf(dl2);
            float x[3] = {c[0] + r[0] * (dir[0] * inv), c[1] + r[0] * (dir[1] * inv), c[2] + r[0] * (dir[2] * inv)};
            dd = (x[0] - pn[0]) * (x[0] - pn[0]) + (x[1] - pn[1]) * (x[1] - pn[1]) + (x[2] - pn[2]) * (x[2] - pn[2]);
        }
        return wyvill_sq(dd);
    }
    case PSGPU_PRIM_LINE: { /* :1042-1048 + NearestPointInLineSegment FW/PS_GeometryFuncs.cpp:156-179 */
        float ld[3] = {d[0] - c[0], d[1] - c[1], d[2] - c[2]}, np[3];
        if (ld[0] == 0.0f && ld[1] == 0.0f && ld[2] == 0.0f) { np[0] = c[0]; np[1] = c[1]; np[2] = c[2]; }
        else {
            float delta = ((pn[0] - c[0]) * ld[0] + (pn[1] - c[1]) * ld[1] + (pn[2] - c[2]) * ld[2]) /
                          (ld[0] * ld[0] + ld[1] * ld[1] + ld[2] * ld[2]);
            if (delta < 0) delta = 0; else if (delta > 1) delta = 1;
            np[0] = c[0] + delta * ld[0]; np[1] = c[1] + delta * ld[1]; np[2] = c[2] + delta * ld[2];
        }
        float dx = np[0] - pn[0], dy = np[1] - pn[1], dz = np[2] - pn[2];
        return wyvill_sq(dx * dx + dy * dy + dz * dz);
    }
    case PSGPU_PRIM_QUADRICPOINT: { /* :1051-1064 (BT/CQuadricPoint.h:43-52): scale * (1 - d^2 / R^2)^2 inside R */
        float dx = pn[0] - c[0], dy = pn[1] - c[1], dz = pn[2] - c[2];
        float fDist2 = dx * dx + dy * dy + dz * dz;
        float fValue = 1.0f - (fDist2 / (r[0] * r[0]));
        if (fValue <= 0.0f) return 0.0f;
        return r[1] * fValue * fValue;
    }
    case PSGPU_PRIM_INSTANCE: { /* :1069-1078: the origin node at the instance's local point, w = 1, no stores */
        int origin = (int)r[0];
        if (origin < 0 || origin >= s->n || origin == id) return 0.0f;
        return scalar_node(s, origin, pn, 1.0f, NULL);
    }
    default: return 0.0f; /* bntPrimNull :1066-1068 */
    }
}

/* fieldvalueOp, PH/CompactBlobTree.cpp:677-891: own affine -> warp -> kids -> combine. `store`
 * (optional, n floats) receives every node's field like lpStoreFVOp / lpStoreFVPrim. */
/* `w` is the 4th lane of the reference's vec4 point: 0 from fieldvalue() (:478-479) and after any warp
 * (Vector4d() zero-inits w, FW/PS_Vector.h:1004-1007), 1 right after an operator affine (:693); an
 * operator's own matrix therefore drops its translation unless the parent was an affine-only operator. */
static float scalar_node(const pso_scalar* s, int id, const float* p, float w, float* store) {
    int t = s->type[id];
    if (PSGPU_IS_PRIM(t)) {
        float f = scalar_prim(s, id, p);
        if (store) store[id] = f;
        return f;
    }
    float pw[3] = {p[0], p[1], p[2]};
    if (s->hasxf[id]) { /* :685-694 */
        const float* m = &s->xf[id * 12];
        pw[0] = m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3] * w;
        pw[1] = m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7] * w;
        pw[2] = m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11] * w;
        w = 1.0f;
    }
    const float* r = &s->res[id * 4];
    float q[3];
    switch (t) { /* :712-746 */
    case PSGPU_OP_WARP_BEND: warp_bend(pw, r[0], r[1], r[2], r[3], q); memcpy(pw, q, 12); break;
    case PSGPU_OP_WARP_TWIST: warp_twist(pw, r[0], (int)r[1], q); memcpy(pw, q, 12); break;
    case PSGPU_OP_WARP_TAPER: warp_taper(pw, r[0], (int)r[1], (int)r[2], q); memcpy(pw, q, 12); break;
    case PSGPU_OP_WARP_SHEAR: warp_shear(pw, r[0], (int)r[1], (int)r[2], q); memcpy(pw, q, 12); break;
    default: break;
    }
    if (PSGPU_IS_WARP(t)) w = 0.0f;
    int nk = s->nkids[id];
    const int* kids = &s->child[s->first[id]];
    float res = 0.0f;
    for (int i = 0; i < nk; i++) {
        float f = scalar_node(s, kids[i], pw, w, store);
        switch (t) { /* :808-880, folded into the child loop in the same order */
        case PSGPU_OP_BLEND: res += f; break;
        case PSGPU_OP_RICCI: res += powf(f, r[0]); break;
        case PSGPU_OP_UNION: if (i == 0) res = f; else if (f > res) res = f; break;
        case PSGPU_OP_INTERSECT: if (i == 0) res = f; else if (f < res) res = f; break;
        case PSGPU_OP_DIF: if (i == 0) res = f; else res = (res < 1.0f - f) ? res : (1.0f - f); break;
        case PSGPU_OP_SMOOTHDIF: if (i == 0) res = f; else res *= (1.0f - f); break;
        default: if (i == 0) res = f; break; /* warps: first child */
        }
    }
    if (t == PSGPU_OP_RICCI) res = powf(res, r[1]);
    if (store) store[id] = res;
    return res;
}

static float scalar_field(const pso_scalar* s, const float* p, float* store) { return scalar_node(s, s->root, p, 0.0f, store); }

/* baseColorOp with cached node fields, PH/CompactBlobTree.cpp:1124-1294 (evaluated at the un-warped p) */
static void scalar_color(const pso_scalar* s, int id, const float* store, float* out) {
    int t = s->type[id];
    if (t == PSGPU_PRIM_INSTANCE) { /* baseColorPrim :1109-1122: the ORIGIN's colour from the stores of the origin's own evaluation */
        int origin = (int)s->res[id * 4];
        if (origin >= 0 && origin < s->n && origin != id) { scalar_color(s, origin, store, out); return; }
    }
    if (PSGPU_IS_PRIM(t)) { memcpy(out, &s->col[id * 3], 12); return; }
    int nk = s->nkids[id];
    const int* kids = &s->child[s->first[id]];
    out[0] = out[1] = out[2] = 0.0f;
    if (nk == 0) return;
    float acc[3] = {0, 0, 0}, sumF = 0.0f, best = 0.0f, first[3] = {0, 0, 0}, pick[3] = {0, 0, 0};
    for (int i = 0; i < nk; i++) {
        float c[3];
        scalar_color(s, kids[i], store, c);
        float f = store[kids[i]];
        if (i == 0) { memcpy(first, c, 12); memcpy(pick, c, 12); }
        switch (t) {
        case PSGPU_OP_BLEND: case PSGPU_OP_RICCI:
            for (int a = 0; a < 3; a++) acc[a] += c[a] * f;
            sumF += f; break;
        case PSGPU_OP_UNION: if (i == 0) best = f; else if (f > best) { best = f; memcpy(pick, c, 12); } break;
        case PSGPU_OP_INTERSECT: if (i == 0) best = f; else if (f < best) { best = f; memcpy(pick, c, 12); } break;
        case PSGPU_OP_DIF: case PSGPU_OP_SMOOTHDIF:
            if (i == 0) best = f; else { float cf = 1.0f - f; if (cf < best) { best = cf; memcpy(pick, c, 12); } }
            break;
        default: break;
        }
    }
    if (t == PSGPU_OP_BLEND || t == PSGPU_OP_RICCI) {
        if (sumF == 0.0f) memcpy(out, first, 12);
        else { float inv = 1.0f / sumF; for (int a = 0; a < 3; a++) out[a] = acc[a] * inv; }
    } else if (PSGPU_IS_WARP(t)) memcpy(out, first, 12);
    else memcpy(out, pick, 12);
}

/* COMPACTBLOBTREE::normal, PH/CompactBlobTree.cpp:433-450 (exact normalise) */
static void scalar_normal(const pso_scalar* s, const float* p, float f, float delta, float* n) {
    float q[3], a[3];
    for (int i = 0; i < 3; i++) { memcpy(q, p, 12); q[i] = p[i] + delta; a[i] = scalar_field(s, q, NULL); }
    float inv = -1.0f / delta;
    for (int i = 0; i < 3; i++) a[i] = (a[i] - f) * inv;
    float len = sqrtf(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
    if (len > 0.0f) { float il = 1.0f / len; n[0] = a[0] * il; n[1] = a[1] * il; n[2] = a[2] * il; }
    else { n[0] = n[1] = n[2] = 1.0f; } /* normalizeXYZ, FW/PS_Vector.h:1072-1086 */
}

/* ComputeRootNewtonRaphsonVEC4, PH/CompactBlobTree.cpp:1581-1622 */
static int scalar_newton(const pso_scalar* s, const float* p1, const float* p2, float fp1, float fp2, float target,
                         int iterations, float* out, float* outF, float* store) {
    float x[3];
    if (iterations <= 0) return -1;
    if (fabsf(fp1 - target) < fabsf(fp2 - target)) memcpy(x, p1, 12); else memcpy(x, p2, 12);
    int i;
    memcpy(out, x, 12); *outF = 0.0f;
    for (i = 0; i < iterations; i++) {
        /* fieldValueAndGradient :452-474, delta = FIELD_VALUE_EPSILON */
        float fp = scalar_field(s, x, NULL), g[3], q[3];
        float invd = 1.0f / PSO_FIELDVALUE_EPSILON;
        for (int a = 0; a < 3; a++) { memcpy(q, x, 12); q[a] = x[a] + PSO_FIELDVALUE_EPSILON; g[a] = scalar_field(s, q, NULL); }
        for (int a = 0; a < 3; a++) g[a] = (g[a] - fp) * invd;
        float d = target - fp;
        /* grad is a vec4 whose w lane holds fp (:472) and Vector4d::dot sums all four lanes
         * (FW/PS_Vector.h:1090-1093), so the reference's step is damped by fp*fp. Kept as is. */
        float gg = 1.0f / (g[0] * g[0] + g[1] * g[1] + g[2] * g[2] + fp * fp);
        for (int a = 0; a < 3; a++) x[a] = x[a] + (d * g[a]) * gg;
        *outF = scalar_field(s, x, store);
        memcpy(out, x, 12);
        if (fabsf(*outF - target) < PSO_FIELDVALUE_EPSILON) break;
    }
    return (i + 1) * 4;
}

void pso_scalar_field(const pso_scalar* s, int64_t n, const float* xyz, float* out) {
    for (int64_t i = 0; i < n; i++) out[i] = scalar_field(s, &xyz[i * 3], NULL);
}
void pso_scalar_field_color_normal(const pso_scalar* s, int64_t n, const float* xyz, float* f, float* col, float* nrm) {
    float* store = (float*)calloc((size_t)s->n, sizeof(float));
    for (int64_t i = 0; i < n; i++) {
        f[i] = scalar_field(s, &xyz[i * 3], store);
        scalar_color(s, s->root, store, &col[i * 3]);
        scalar_normal(s, &xyz[i * 3], f[i], PSO_NORMAL_DELTA, &nrm[i * 3]);
    }
    free(store);
}
void pso_scalar_newton(const pso_scalar* s, int64_t n, const float* p1, const float* p2, const float* f1,
                       const float* f2, float iso, int iterations, float* root, float* froot, int32_t* evals) {
    for (int64_t i = 0; i < n; i++)
        evals[i] = scalar_newton(s, &p1[i * 3], &p2[i * 3], f1[i], f2[i], iso, iterations, &root[i * 3], &froot[i], NULL);
}

/* doMarchingCubes, PH/CPolyParsipOptimized.cpp:130-327: eager cache here (the lazy FLT_MIN cache only
 * skips corners of cells that are never visited - every corner of a 7^3 block is visited), inside = f > iso,
 * Newton root (DEFAULT_ITERATIONS 8, BT/_constSettings.h:8), normal, baseColor. */
static void scalar_fill_cache(const pso_scalar* s, const float* lo, float cellsize, float* fv) {
    for (int i = 0; i < GD; i++) for (int j = 0; j < GD; j++) for (int k = 0; k < GD; k++) {
        float p[3] = {lo[0] + cellsize * i, lo[1] + cellsize * j, lo[2] + cellsize * k};
        fv[i * GD + j * GD * GD + k] = scalar_field(s, p, NULL);
    }
}
static int scalar_cell_case(const float* fv, int i, int j, int k) {
    int ip0 = i * GD, ip1 = ip0 + GD, jp0 = j * GD * GD, jp1 = jp0 + GD * GD;
    float c[8] = {fv[ip0 + jp0 + k], fv[ip0 + jp0 + k + 1], fv[ip0 + jp1 + k], fv[ip0 + jp1 + k + 1],
                  fv[ip1 + jp0 + k], fv[ip1 + jp0 + k + 1], fv[ip1 + jp1 + k], fv[ip1 + jp1 + k + 1]};
    int idx = 0;
    for (int b = 0; b < 8; b++) if (c[b] > PSO_ISO_VALUE) idx |= 1 << b;
    return idx;
}
static void scalar_process_mpu(const pso_scalar* s, float* store, float cellsize, pso_mpu* mpu) {
    float fv[GD * GD * GD];
    mpu->nv = mpu->nt = 0; mpu->fieldEvals = 0;
    if (s->octLo) {
        const float side = (float)(GD - 1) * cellsize;
        const float hi[3] = {mpu->lo[0] + side, mpu->lo[1] + side, mpu->lo[2] + side};
        int hit = 0;
        for (int i = 0; i < s->n && !hit; i++) {
            if (!PSGPU_IS_PRIM(s->type[i])) continue;
            const float *a = &s->octLo[i * 3], *b = &s->octHi[i * 3];
            hit = !((a[0] >= hi[0]) || (b[0] <= mpu->lo[0]) || (a[1] >= hi[1]) || (b[1] <= mpu->lo[1]) || (a[2] >= hi[2]) || (b[2] <= mpu->lo[2]));
        }
        if (!hit) return;
    }
    scalar_fill_cache(s, mpu->lo, cellsize, fv);
    mpu->fieldEvals = GD * GD * GD;
    unsigned char hasEdge[GD * GD * GD * 3];
    uint16_t edgeVid[GD * GD * GD * 3];
    memset(hasEdge, 0, sizeof(hasEdge));
    for (int i = 0; i < GD - 1; i++) for (int j = 0; j < GD - 1; j++) for (int k = 0; k < GD - 1; k++) {
        int cfg = scalar_cell_case(fv, i, j, k);
        if (cfg == 0 || cfg == 255) continue;
        const int* row = &kTriTable[cfg * 16];
        int vids[16], nEdges = 0;
        for (int ic = 0; ic < 16 && row[ic] != -1; ic++) {
            int cand = row[ic];
            int c1 = kCorner1[cand], c2 = kCorner2[cand], ax = kEdgeAxis[cand];
            int s1[3] = {i + ((c1 & 4) >> 2), j + ((c1 & 2) >> 1), k + (c1 & 1)};
            int s2[3] = {i + ((c2 & 4) >> 2), j + ((c2 & 2) >> 1), k + (c2 & 1)};
            int h = s1[0] * GD + s1[1] * GD * GD + s1[2];
            if (hasEdge[h * 3 + ax]) vids[ic] = edgeVid[h * 3 + ax];
            else {
                float p1[3] = {mpu->lo[0] + cellsize * s1[0], mpu->lo[1] + cellsize * s1[1], mpu->lo[2] + cellsize * s1[2]};
                float p2[3] = {mpu->lo[0] + cellsize * s2[0], mpu->lo[1] + cellsize * s2[1], mpu->lo[2] + cellsize * s2[2]};
                float f1 = fv[s1[0] * GD + s1[1] * GD * GD + s1[2]], f2 = fv[s2[0] * GD + s2[1] * GD * GD + s2[2]];
                float p[3], fp;
                int v = mpu->nv;
                mpu->fieldEvals += (uint32_t)scalar_newton(s, p1, p2, f1, f2, PSO_ISO_VALUE, 8, p, &fp, store);
                mpu->fieldEvals += 3;
                scalar_normal(s, p, fp, PSO_NORMAL_DELTA, &mpu->nrm[v * 3]);
                scalar_color(s, s->root, store, &mpu->col[v * 3]);
                memcpy(&mpu->pos[v * 3], p, 12);
                mpu->nv++;
                vids[ic] = v; hasEdge[h * 3 + ax] = 1; edgeVid[h * 3 + ax] = (uint16_t)v;
            }
            nEdges++;
        }
        for (int t = 0; t < nEdges / 3; t++) {
            mpu->tri[mpu->nt * 3 + 0] = (uint16_t)vids[t * 3]; mpu->tri[mpu->nt * 3 + 1] = (uint16_t)vids[t * 3 + 1];
            mpu->tri[mpu->nt * 3 + 2] = (uint16_t)vids[t * 3 + 2];
            mpu->nt++;
        }
    }
}
static void scalar_mpu_tramp(const void* sc, void* scratch, float cellsize, pso_mpu* mpu) {
    scalar_process_mpu((const pso_scalar*)sc, (float*)scratch, cellsize, mpu);
}
void pso_scalar_mpu_cache(const pso_scalar* s, const float* lo0, float cellsize, const int32_t* ijk, float* fv512, uint8_t* case343) {
    float side = (float)(GD - 1) * cellsize;
    float lo[3] = {lo0[0] + side * (float)ijk[0], lo0[1] + side * (float)ijk[1], lo0[2] + side * (float)ijk[2]};
    scalar_fill_cache(s, lo, cellsize, fv512);
    for (int i = 0; i < GD - 1; i++) for (int j = 0; j < GD - 1; j++) for (int k = 0; k < GD - 1; k++)
        case343[i * 49 + j * 7 + k] = (uint8_t)scalar_cell_case(fv512, i, j, k);
}
int pso_scalar_polygonize(pso_scalar* s, const float* lo, const float* hi, float cellsize, int nthreads,
                          const int32_t* b0, const int32_t* e0, int keep_mesh, int64_t* stats) {
    memcpy(s->lo, lo, 12); memcpy(s->hi, hi, 12);
    return run_driver(s, 0, scalar_mpu_tramp, cellsize, s->lo, s->hi, nthreads, b0, e0, keep_mesh, &s->mesh, stats);
}
void pso_scalar_result_counts(const pso_scalar* s, int64_t* o) { o[0] = s->mesh.nv; o[1] = s->mesh.nt; o[2] = s->mesh.nm; }
void pso_scalar_result_copy(const pso_scalar* s, float* pos, float* nrm, float* col, uint32_t* tri, int32_t* mi, uint32_t* mc) {
    result_copy(&s->mesh, pos, nrm, col, tri, mi, mc);
}

/* ---------------------------------------------------------------------------------------------
 * driver: static lattice, dynamic chunks, results merged in lattice order so the output does not
 * depend on the thread count
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    drv_ctx* c; int tid;
    kept_mpu* kept; int64_t nkept, capkept;
    int64_t nproc, ncross, nv, nt, fe, nover;
} drv_thread;

static void* drv_worker(void* arg) {
    drv_thread* t = (drv_thread*)arg;
    drv_ctx* c = t->c;
    void* scratch;
    simd_scratch sw;
    if (c->isSimd) { scratch_init((const pso_simd*)c->scene, &sw); scratch = &sw; }
    else scratch = calloc((size_t)((const pso_scalar*)c->scene)->n, sizeof(float));
    pso_mpu* mpu = (pso_mpu*)malloc(sizeof(pso_mpu));
    const int ny = c->E[1] - c->B[1], nz = c->E[2] - c->B[2];
    const float side = c->cellsize * (float)(GD - 1);
    const int64_t CH = 16;
    for (;;) {
        int64_t b = __sync_fetch_and_add(c->next, CH);
        if (b >= c->total) break;
        int64_t e = b + CH < c->total ? b + CH : c->total;
        for (int64_t q = b; q < e; q++) {
            int i = c->B[0] + (int)(q / ((int64_t)ny * nz)), j = c->B[1] + (int)((q / nz) % ny), k = c->B[2] + (int)(q % nz);
            mpu->lo[0] = c->lo[0] + (float)i * side; mpu->lo[1] = c->lo[1] + (float)j * side; mpu->lo[2] = c->lo[2] + (float)k * side;
            c->fn(c->scene, scratch, c->cellsize, mpu);
            t->nproc++; t->fe += mpu->fieldEvals; t->nv += mpu->nv; t->nt += mpu->nt;
            if (mpu->nv > 512 || mpu->nt > 512) t->nover++;
            if (mpu->nt > 0) {
                t->ncross++;
                if (c->keep) {
                    if (t->nkept == t->capkept) { t->capkept = t->capkept * 2 + 64; t->kept = (kept_mpu*)realloc(t->kept, sizeof(kept_mpu) * (size_t)t->capkept); }
                    /* compact copy */
                    pso_mpu* cp = (pso_mpu*)malloc(sizeof(pso_mpu));
                    memcpy(cp, mpu, sizeof(pso_mpu));
                    t->kept[t->nkept].key = q; t->kept[t->nkept].mpu = cp; t->nkept++;
                }
            }
        }
    }
    free(mpu);
    if (c->isSimd) scratch_free(&sw); else free(scratch);
    return NULL;
}

static int cmp_kept(const void* a, const void* b) {
    int64_t x = ((const kept_mpu*)a)->key, y = ((const kept_mpu*)b)->key;
    return x < y ? -1 : (x > y ? 1 : 0);
}

static int run_driver(const void* scene, int isSimd, mpu_fn fn, float cellsize, const float* lo, const float* hi,
                      int nthreads, const int32_t* b0, const int32_t* e0, int keep, pso_mesh* mesh, int64_t* stats) {
    int nm[3];
    lattice_dims(cellsize, lo, hi, nm);
    drv_ctx c;
    memset(&c, 0, sizeof(c));
    c.scene = scene; c.fn = fn; c.cellsize = cellsize; c.lo = lo; c.keep = keep; c.isSimd = isSimd;
    for (int a = 0; a < 3; a++) {
        c.B[a] = b0 ? b0[a] : 0; c.E[a] = e0 ? e0[a] : nm[a];
        if (c.B[a] < 0) c.B[a] = 0;
        if (c.E[a] > nm[a]) c.E[a] = nm[a];
        if (c.E[a] < c.B[a]) c.E[a] = c.B[a];
    }
    c.total = (int64_t)(c.E[0] - c.B[0]) * (c.E[1] - c.B[1]) * (c.E[2] - c.B[2]);
    volatile int64_t next = 0;
    c.next = &next;
    if (nthreads < 1) nthreads = 1;
    drv_thread* th = (drv_thread*)calloc((size_t)nthreads, sizeof(drv_thread));
    pthread_t* pt = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    pthread_attr_t attr; pthread_attr_init(&attr); pthread_attr_setstacksize(&attr, (size_t)64 << 20);
    for (int t = 0; t < nthreads; t++) { th[t].c = &c; th[t].tid = t; pthread_create(&pt[t], &attr, drv_worker, &th[t]); }
    for (int t = 0; t < nthreads; t++) pthread_join(pt[t], NULL);
    pthread_attr_destroy(&attr);
    int64_t tot[6] = {0, 0, 0, 0, 0, 0}, nk = 0;
    for (int t = 0; t < nthreads; t++) {
        tot[0] += th[t].nproc; tot[1] += th[t].ncross; tot[2] += th[t].nv; tot[3] += th[t].nt; tot[4] += th[t].fe; tot[5] += th[t].nover;
        nk += th[t].nkept;
    }
    mesh_clear(mesh);
    if (keep) {
        kept_mpu* all = (kept_mpu*)malloc(sizeof(kept_mpu) * (size_t)(nk ? nk : 1));
        int64_t o = 0;
        for (int t = 0; t < nthreads; t++) { memcpy(all + o, th[t].kept, sizeof(kept_mpu) * (size_t)th[t].nkept); o += th[t].nkept; }
        qsort(all, (size_t)nk, sizeof(kept_mpu), cmp_kept);
        const int ny = c.E[1] - c.B[1], nz = c.E[2] - c.B[2];
        for (int64_t q = 0; q < nk; q++) {
            int64_t key = all[q].key;
            int ijk[3] = {c.B[0] + (int)(key / ((int64_t)ny * nz)), c.B[1] + (int)((key / nz) % ny), c.B[2] + (int)(key % nz)};
            mesh_append(mesh, all[q].mpu, ijk);
            free(all[q].mpu);
        }
        free(all);
    }
    for (int t = 0; t < nthreads; t++) free(th[t].kept);
    free(th); free(pt);
    if (stats) {
        stats[0] = tot[0]; stats[1] = tot[1]; stats[2] = tot[2]; stats[3] = tot[3]; stats[4] = tot[4];
        stats[5] = nm[0]; stats[6] = nm[1]; stats[7] = nm[2]; stats[8] = tot[5];
    }
    return PSGPU_SUCCESS;
}
