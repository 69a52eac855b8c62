// Postfix instruction stream of a linearized BlobTree, shared by the host linearizer and the kernels.
//
// The stream is a sequence of 16-byte "quads" (4 x 32-bit words) so that every parameter block is
// one or two LDS.128 / LDG.128 and a whole program is a legal cp.async.bulk (TMA) transfer.
// Quad 0 of an instruction is {header, p0, p1, p2}; the header word is
//     bits 0..4   opcode
//     bits 5..8   combine code (how a primitive's field, or a popped sub-tree value, meets the stack top)
//     bit  9      XF: three more quads (rows of a backward 3x4 matrix) follow the primitive's own quads
//     bit  10     PRE: the colour-scale count below applies to THIS primitive's colour before it is combined (left by the culling
//                 pass when it folds a one-primitive group "p:PUSH ... OPC_COMBINE" into "p:<combine>")
//     bits 11..26 primitive id (index into the colour table)
//     bits 27..31 colour-scale count: after this instruction the colour walk multiplies the top colour that many times by
//                 (2*(0.5+top)-1) - the weight a PS_SimdPoly blend gives its surviving operand when the culling pass dropped
//                 the all-zero other one (see OPC_CSCALE, which carries counts beyond 31). Free for the field-only passes.
//
// Evaluation model: a value stack whose top lives in registers, and a point stack for warps. A primitive
// evaluates the Wyvill field (1-d^2)^3 of its skeleton distance at the current point and combines it into
// the stack top (or pushes). n-ary operators therefore cost no instruction of their own for primitive
// children: blend(a, b, c) = a:PUSH b:ADD c:ADD. Sub-tree children are evaluated (PUSH ...) and folded in
// with OPC_COMBINE. Warps bracket their sub-tree: OPC_TWIST ... OPC_POPPOINT.
//
// Reference semantics restated: primitives PS_SimdPoly/include/PS_Polygonizer.cpp:934-1179 and
// ParsipHaptics/include/CompactBlobTree.cpp:893-1092; operators PS_Polygonizer.cpp:1281-1338 and
// CompactBlobTree.cpp:808-880; warps CompactBlobTree.cpp:1315-1536.
#ifndef PSGPU_PROGRAM_H
#define PSGPU_PROGRAM_H
#include <stdint.h>

enum PsOpcode : uint32_t {
    OPC_END = 0,
    // primitives (opcode < OPC_FIRST_NONPRIM)
    OPC_POINT = 1,     // {h, px,py,pz}
    OPC_SEGMENT = 2,   // {h, ax,ay,az}{dx,dy,dz, 1/|d|^2}    clamped (scalar twin)
    OPC_LINE_INF = 3,  // {h, ax,ay,az}{dx,dy,dz, |d|^2}      unclamped (SIMD path)
    OPC_CYLINDER = 4,  // {h, px,py,pz}{dx,dy,dz, radius}{height,0,0,0}
    OPC_CUBE = 5,      // {h, px,py,pz}{side,0,0,0}
    OPC_DISC = 6,      // {h, px,py,pz}{nx,ny,nz, radius}
    OPC_RING = 7,      // {h, px,py,pz}{nx,ny,nz, radius}
    OPC_NULLPRIM = 8,  // {h,0,0,0}  field 0
    OPC_QUADRIC = 9,   // {h, px,py,pz}{1/R^2, scale, R, 0}  quadric point: scale * (1 - d^2/R^2)^2 inside R (CQuadricPoint.h:43-52);
                       //                                    native semantics only, the only primitive whose law is not Wyvill's
    OPC_FIRST_NONPRIM = 16,
    OPC_COMBINE = 16,    // {h}            v = pop; top = top (combine) v
    OPC_POW = 17,        // {h, e}         top = pow(top, e)          (true Ricci, CompactBlobTree.cpp:813-818); combine code 1 marks
                         //                the per-child power (the colour walk remembers the raw value), 0 the closing 1/n power
    OPC_RICCI_SIMD = 18, // {h, e}         v = pop; top = fast_pow(top + v, e)   (PS_SIMDVecN.h:122-128, Q6)
    OPC_CLAMP1 = 19,     // {h}            top = min(top, 1)          (difference with an all-zero subtrahend, Q16)
    OPC_DIF0 = 20,       // {h}            top = min(0, 1 - top)      (difference with an all-zero minuend)
    OPC_CSCALE = 21,     // {h, n}         colour *= (2*(0.5+top)-1), n times; field untouched. Left by the culling pass where a
                         //                blend lost an all-zero operand: the PS_SimdPoly colour rule 2*f_L*c_L + 2*f_R*c_R is
                         //                unnormalised (PS_Polygonizer.cpp:1475-1487), so the surviving operand's colour keeps its weight
    OPC_BOXMASK = 22,    // {h}{lo.xyz, 0}{hi.xyz, 0}   PS_SimdPoly semantics linearized with PSGPU_LIN_AS_SHIPPED only: top = 0 for a PACKET of
                         //                points of which none lies inside the box's x-slab OR y-slab OR z-slab - the reference's prune of
                         //                operators deeper than 3 levels (PS_Polygonizer.cpp:1228-1252). A packet = the points
                         //                the reference evaluates in one SIMD call: 4 z-consecutive corners in the grid phase, the 4 root
                         //                samples of an edge, a single point everywhere else
    OPC_TWIST = 24,      // {h, k, axis}                 push point; warp
    OPC_TAPER = 25,      // {h, k, along, taper}
    OPC_BEND = 26,       // {h, k, y0, left}{right, 1/k, 0, 0}
    OPC_SHEAR = 27,      // {h, k, along, dep}
    OPC_AFFINE = 28,     // {h,0,0,0}{r0}{r1}{r2}        push point; p = M p (operator-level matrix, Q14)
    OPC_POPPOINT = 29,   // {h}                          pop point
};

enum PsCombine : uint32_t {
    CMB_PUSH = 0,
    CMB_ADD = 1,   // blend
    CMB_MAX = 2,   // union
    CMB_MIN = 3,   // intersection
    CMB_DIF = 4,   // top = min(top, 1 - v)
    CMB_SDIF = 5,  // top = top * (1 - v)
    CMB_RDIF = 7,  // top = min(v, 1 - top)   the incoming value is the MINUEND (left operand) ...
    CMB_RSDIF = 8, // top = v * (1 - top)     ... of a (smooth) difference whose right operand was evaluated first: lets the
                   // linearizer walk the right-leaning chains ConvertToBinaryTree produces (op(prim, op(prim, ...))) innermost
                   // first with ONE live stack entry, bit-identical to the reference's association (a + b == b + a in IEEE)
    CMB_RADD = 6,  // top + v like CMB_ADD; sums the POWERED children of a true Ricci blend, whose colour weights are the RAW
                   // child fields (CompactBlobTree.cpp:1186-1203), kept in the colour walk's aux slot
};

#define PS_HDR_CMB_SHIFT 5
#define PS_HDR_CSC_SHIFT 27
#define PS_HDR_CSC_MAX 31u
#define PS_HDR_PRE_BIT (1u << 10)
#define PS_HDR(opc, cmb, xf, prim) ((uint32_t)(opc) | ((uint32_t)(cmb) << PS_HDR_CMB_SHIFT) | ((uint32_t)((xf) ? 1 : 0) << 9) | ((uint32_t)(prim) << 11))
#define PS_HDR_OPC(h) ((h) & 31u)
#define PS_HDR_CMB(h) (((h) >> PS_HDR_CMB_SHIFT) & 15u)
#define PS_HDR_XF(h) (((h) >> 9) & 1u)
#define PS_HDR_PRE(h) (((h) >> 10) & 1u)
#define PS_HDR_PRIM(h) (((h) >> 11) & 0xFFFFu)
#define PS_HDR_CSC(h) ((h) >> PS_HDR_CSC_SHIFT)
#define PS_HDR_SET_CMB(h, cmb) (((h) & ~(15u << PS_HDR_CMB_SHIFT)) | ((uint32_t)(cmb) << PS_HDR_CMB_SHIFT))
#define PS_MAX_PRIMS (1u << 16)

// quads per instruction, without the optional XF rows
#if defined(__CUDACC__)
__host__ __device__
#endif
static inline int ps_base_quads(uint32_t opc) {
    // 2 bits per opcode (quads - 1), opcodes 0..15 in the first word and 16..31 in the second: a shift and a mask instead of a
    // compare chain in the interpreters' decode path. SEGMENT, LINE_INF, CUBE, DISC, RING, BEND = 2; CYLINDER, BOXMASK = 3; AFFINE = 4.
    const uint32_t k = (opc < 16u) ? 0x00045650u : 0x03102000u;  // QUADRIC (9) = 2; BOXMASK (22) = 3
    return 1 + (int)((k >> (2u * (opc & 15u))) & 3u);
}
#if defined(__CUDACC__)
__host__ __device__
#endif
static inline int ps_instr_quads(uint32_t hdr) { return ps_base_quads(PS_HDR_OPC(hdr)) + (PS_HDR_XF(hdr) ? 3 : 0); }

// SURVEY.md 8(d) algorithmic FP32 op counts (FMA = 2), incl. the 4-op Wyvill tail; +18 with a matrix.
#if defined(__CUDACC__)
__host__ __device__
#endif
static inline int ps_algorithmic_flops(uint32_t hdr) {
    int f = 0;
    switch (PS_HDR_OPC(hdr)) {
    case OPC_POINT: case OPC_QUADRIC: f = 12; break;
    case OPC_SEGMENT: case OPC_LINE_INF: f = 27; break;
    case OPC_CYLINDER: f = 33; break;
    case OPC_CUBE: f = 39; break;
    case OPC_DISC: case OPC_RING: f = 51; break;
    case OPC_COMBINE: f = (PS_HDR_CMB(hdr) == CMB_DIF || PS_HDR_CMB(hdr) == CMB_SDIF) ? 2 : 1; break;
    case OPC_POW: f = 2; break;
    case OPC_RICCI_SIMD: f = 4; break;
    case OPC_TWIST: f = 9; break;
    case OPC_TAPER: f = 3; break;
    case OPC_BEND: f = 16; break;
    case OPC_SHEAR: f = 2; break;
    case OPC_AFFINE: f = 18; break;
    default: f = 0; break;
    }
    if (PS_HDR_OPC(hdr) < OPC_FIRST_NONPRIM && PS_HDR_OPC(hdr) != OPC_END) {
        if (PS_HDR_XF(hdr)) f += 18;
        if (PS_HDR_CMB(hdr) != CMB_PUSH) f += (PS_HDR_CMB(hdr) == CMB_DIF || PS_HDR_CMB(hdr) == CMB_SDIF || PS_HDR_CMB(hdr) >= CMB_RDIF) ? 2 : 1;
    }
    return f;
}

#define PS_MAX_VALUE_STACK 32
#define PS_MAX_POINT_STACK 8

#define PS_GRID_DIM 8
#define PS_CELLS_PER_MPU 7
#define PS_ISO_VALUE 0.5f
#define PS_NORMAL_DELTA 0.001f
#define PS_FIELDVALUE_EPSILON 0.001f

#endif
