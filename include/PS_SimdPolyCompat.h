// Drop-in C++ facade for the reference's PS_SimdPoly entry points, backed by libpsgpu.so.
//
// Include this INSTEAD of Parsip100/PS_SimdPoly/include/PS_Polygonizer.h: same namespace, type names,
// field names, array capacities, return codes and call signatures for the polygonization path
// (PS_Polygonizer.h:47-50, 70-81, 96-198, 384-393), so ParsipCmd / PS_FrameWork callers recompile
// unchanged. The structs are layout-compatible with the reference's (sizes pinned by
// the static_asserts below: sizeof(MPU) == 21524, sizeof(PolyMPUs) == 516576004).
//
// What is different by design: the work runs on the GPU through the C ABI of psgpu.h (no TBB, no CPU
// fallback); Polygonize returns RET_NOT_ENOUGH_MEM (-2) where the reference silently overflows an MPU's
// 512-vertex / 512-triangle arrays; MPUSTATS carries no TBB types.
#ifndef PS_SIMDPOLY_COMPAT_H
#define PS_SIMDPOLY_COMPAT_H
#include <stdint.h>

typedef unsigned char U8;
typedef unsigned short U16;
typedef unsigned int U32;

namespace PS {
namespace MATHFUNCTIONAL {
template <typename T>
struct svec3 {
    svec3() {}
    explicit svec3(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
    T x, y, z;
};
typedef svec3<float> svec3f;
typedef svec3<int> svec3i;
}  // namespace MATHFUNCTIONAL

namespace SIMDPOLY {
using MATHFUNCTIONAL::svec3f;
using MATHFUNCTIONAL::svec3i;

#define GRID_DIM 8
#define RET_PARAM_ERROR -1
#define RET_NOT_ENOUGH_MEM -2
#define RET_INVALID_BVH -3
#define RET_SUCCESS 1

#define ISO_VALUE 0.5f
#define ISO_DIST 0.45420206f
#define FIELDVALUE_EPSILON 0.001f
#define NORMAL_DELTA 0.001f
#define MIN_CELL_SIZE 0.01f
#define MAX_CELL_SIZE 0.20f
#define DEFAULT_CELL_SIZE 0.14f

#define MAX_TREE_NODES 128
#define MAX_MATRIX_COUNT MAX_TREE_NODES
#define PRIM_MATRIX_STRIDE 12
#define BOX_MATRIX_STRIDE 16
#define MAX_MPU_COUNT 24000
#define MAX_MPU_VERTEX_COUNT 512
#define MAX_MPU_TRIANGLE_COUNT 512

#ifndef BLOBTREE_DEFINITIONS
// standalone numbering (PS_Polygonizer.h:86-90); with the scene API included first the numbering of
// PS_BlobTree/include/_constSettings.h:26-38 is in force instead - define
// PSGPU_COMPAT_SCENE_API_ENUM before including this header in that case.
enum BlobNodeType { bntPrimCylinder, bntPrimDisc, bntPrimLine, bntPrimPoint, bntPrimRing, bntPrimPolygon, bntPrimCube, bntPrimTriangle,
                    bntPrimCatmullRomCurve, bntPrimSkeleton, bntPrimQuadricPoint, bntPrimFastQuadraticPointSet, bntPrimHalfPlane, bntPrimNull,
                    bntOpUnion, bntOpIntersect, bntOpDif, bntOpSmoothDif, bntOpBlend, bntOpRicciBlend, bntOpGradientBlend, bntOpAffine,
                    bntOpWarpTwist, bntOpWarpTaper, bntOpWarpBend, bntOpWarpShear, bntOpCache, bntOpTexture, bntOpPCM };
#endif

struct SOABlobPrims {
    float posX[MAX_TREE_NODES], posY[MAX_TREE_NODES], posZ[MAX_TREE_NODES];
    float dirX[MAX_TREE_NODES], dirY[MAX_TREE_NODES], dirZ[MAX_TREE_NODES];
    float resX[MAX_TREE_NODES], resY[MAX_TREE_NODES], resZ[MAX_TREE_NODES];
    float colorX[MAX_TREE_NODES], colorY[MAX_TREE_NODES], colorZ[MAX_TREE_NODES];
    float vPrimBoxLoX[MAX_TREE_NODES], vPrimBoxLoY[MAX_TREE_NODES], vPrimBoxLoZ[MAX_TREE_NODES];
    float vPrimBoxHiX[MAX_TREE_NODES], vPrimBoxHiY[MAX_TREE_NODES], vPrimBoxHiZ[MAX_TREE_NODES];
    U8 skeletType[MAX_TREE_NODES];
    U8 idxMatrix[MAX_TREE_NODES];
    svec3f bboxLo;
    svec3f bboxHi;
    U32 ctPrims;
} __attribute__((aligned(4)));

struct SOABlobOps {
    U8 opType[MAX_TREE_NODES];
    U8 opLeftChild[MAX_TREE_NODES];
    U8 opRightChild[MAX_TREE_NODES];
    U8 opChildKind[MAX_TREE_NODES];
    float vBoxLoX[MAX_TREE_NODES], vBoxLoY[MAX_TREE_NODES], vBoxLoZ[MAX_TREE_NODES];
    float vBoxHiX[MAX_TREE_NODES], vBoxHiY[MAX_TREE_NODES], vBoxHiZ[MAX_TREE_NODES];
    float resX[MAX_TREE_NODES], resY[MAX_TREE_NODES], resZ[MAX_TREE_NODES], resW[MAX_TREE_NODES];
    U32 ctOps;
} __attribute__((aligned(4)));

struct SOABlobPrimMatrices {
    float matrix[MAX_MATRIX_COUNT * PRIM_MATRIX_STRIDE];
    U32 count;
} __attribute__((aligned(4)));

struct SOABlobBoxMatrices {
    float matrix[MAX_MATRIX_COUNT * BOX_MATRIX_STRIDE];
    U32 count;
} __attribute__((aligned(4)));

struct MPU {
    float vPos[MAX_MPU_VERTEX_COUNT * 3];
    float vNorm[MAX_MPU_VERTEX_COUNT * 3];
    float vColor[MAX_MPU_VERTEX_COUNT * 3];
    U16 triangles[MAX_MPU_TRIANGLE_COUNT * 3];
    U16 ctVertices;
    U16 ctTriangles;
    svec3f bboxLo;
    U32 ctFieldEvals;
};

struct PolyMPUs {
    MPU vMPUs[MAX_MPU_COUNT];
    U32 ctMPUs;
};

// per-MPU statistics; the reference's version holds TBB thread ids / tick_counts (PS_Polygonizer.h:201-207)
struct MPUSTATS {
    int idxThread;     // GPU: rank of the MPU among evaluated MPUs, -1 if culled
    int bIntersected;  // produced triangles
    double msStart, msEnd;  // whole-run GPU interval (per-MPU timing does not exist on the device)
};

static_assert(sizeof(SOABlobPrims) == 9500, "SOABlobPrims layout");
static_assert(sizeof(SOABlobOps) == 5636, "SOABlobOps layout");
static_assert(sizeof(SOABlobPrimMatrices) == 6148, "SOABlobPrimMatrices layout");
static_assert(sizeof(SOABlobBoxMatrices) == 8196, "SOABlobBoxMatrices layout");
static_assert(sizeof(MPU) == 21524, "MPU layout");
static_assert(sizeof(PolyMPUs) == 516576004ull, "PolyMPUs layout");

U32 CountMPUNeeded(float cellsize, const svec3f& lo, const svec3f& hi);
int PrepareBBoxes(float cellsize, SOABlobPrims& blobPrims, SOABlobBoxMatrices& boxMatrices, SOABlobOps& blobOps);
int Polygonize(float cellsize, const SOABlobPrims& blobPrims, const SOABlobPrimMatrices& blobPrimMatrices, const SOABlobOps& blobOps,
               PolyMPUs& polyMPUs, MPUSTATS* lpProcessStats = 0);
void PrintThreadResults(int ctAttempts, U32* lpThreadProcessed = 0, U32* lpThreadCrossed = 0);

// FieldComputer::fieldValue for arbitrary points (PS_Polygonizer.h:290-291), n points at once
int FieldValues(const SOABlobPrims& blobPrims, const SOABlobPrimMatrices& blobPrimMatrices, const SOABlobOps& blobOps, const float* xyz, int n,
                float* outField);

// FieldComputer::computeRootNewtonRaphson (PS_Polygonizer.h:311-314, .cpp:1625-1681) for n point pairs at once: p1 / p2 are 3 floats per
// pair, fp1 / fp2 their field values; per pair the root, the field there and the member's return value ((rounds) * 4 field evaluations).
int RootsNewtonRaphson(const SOABlobPrims& blobPrims, const SOABlobPrimMatrices& blobPrimMatrices, const SOABlobOps& blobOps, const float* p1,
                       const float* p2, const float* fp1, const float* fp2, int n, float* outRoot, float* outField, int* outEvals,
                       float target_field, int iterations);

}  // namespace SIMDPOLY
}  // namespace PS

// The wrapper class of ParsipHaptics/include/PS_HighPerformanceRender.h:15-33 minus draw() (no GL here).
// linearizeBlobTree(CBlobNode*) keeps the reference signature (PS_HighPerformanceRender.h:27); it is defined inline in
// include/psgpu_blobtree_walk.h, which a TU includes after the reference's scene API (only such a TU can walk a CBlobNode).
// The flat-tree overload is what it forwards to and what callers without the scene API use.
struct psgpu_tree;
namespace PS { namespace BLOBTREE { class CBlobNode; } }
class SimdPoly {
public:
    SimdPoly();
    ~SimdPoly();
    void reset();
    int linearizeBlobTree(PS::BLOBTREE::CBlobNode* root);
    int linearizeBlobTree(const psgpu_tree* root);
    int run(float cellsize);
    const PS::SIMDPOLY::PolyMPUs& polyMPUs() const { return *m_polyMPUs; }
    // the linearized tables (read-only; the reference keeps them private and draws from them)
    const PS::SIMDPOLY::SOABlobPrims& blobPrims() const { return m_blobPrims; }
    const PS::SIMDPOLY::SOABlobOps& blobOps() const { return m_blobOps; }
    const PS::SIMDPOLY::SOABlobPrimMatrices& blobPrimMatrices() const { return m_blobPrimMatrices; }

private:
    PS::SIMDPOLY::SOABlobPrims m_blobPrims;
    PS::SIMDPOLY::SOABlobOps m_blobOps;
    PS::SIMDPOLY::SOABlobPrimMatrices m_blobPrimMatrices;
    PS::SIMDPOLY::SOABlobBoxMatrices m_blobBoxMatrices;
    PS::SIMDPOLY::PolyMPUs* m_polyMPUs;
};
#endif
