// Internal device-side data layout of libmtscuda (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/mtsgpu.h"

#define MTS_EPSILON        1e-4f   // include/mitsuba/core/constants.h:27-35
#define MTS_SHADOW_EPSILON 1e-3f
#define MTS_PI             3.14159265358979323846f
#define MTS_INV_PI         0.31830988618379067154f

// ---- TriAccel record: same 48-byte layout as include/mitsuba/render/triaccel.h:34-46 --------
// Bit 31 of `k` is ours: set when the owning shape has no BSDF, i.e. is not an occluder for
// shadow rays (shape.h:361, skdtree.h:324-334). Closest-hit masks it off.
struct __align__(16) TriAccelRec {
	uint32_t k; float n_u, n_v, n_d;
	float a_u, a_v, b_nu, b_nv;
	float c_nu, c_nv; uint32_t shapeIndex, primIndex;
};
#define TRI_K_MASK         3u
#define TRI_NONOCCLUDER    0x80000000u

// ---- 4-wide BVH node, one 128-byte line: eight 128-bit loads ------------------------------
// child[i]: 0xFFFFFFFF empty | bit31 leaf: (count-1)<<28 | first sorted triangle | else node index
struct __align__(128) BVH4Node {
	float4 lox, loy, loz, hix, hiy, hiz;
	uint4 child;
	uint4 pad;
};
#define BVH_EMPTY      0xFFFFFFFFu
#define BVH_LEAF       0x80000000u
#define BVH_LEAF_MAX   4

// ---- the node the traversal kernels read: the same four children in 64 bytes = two 256-bit loads ---------------
// ncu on the 128-byte node (profiles/r02_base_c4_trace_*): the L1 data pipe ran at 75 % of its wavefront rate — a
// divergent 128-bit load costs one wavefront per lane, eight of them per node visit — while DRAM sat at 11 %.
// Child boxes are quantised to 8 bits per plane on a per-node grid (origin + q * 2^e per axis), rounded OUTWARDS from
// the padded fp32 boxes, so the slab test only ever accepts more than the fp32 test would.
// Decoding costs one PRMT and one FFMA per plane: the byte goes into bits 8..15 of the mantissa of 1.0f, i.e. the
// float 1 + q / 2^15, and `scale` holds 2^(e + 15), so plane = origin + (m - 1) * scale.
struct __align__(64) BVH4QNode {
	float ox, oy, oz, scaleX;      // words 0-3
	uint32_t child[4];             // words 4-7
	uint32_t qlox, qloy, qloz;     // words 8-10: byte i = child i
	uint32_t qhix, qhiy, qhiz;     // words 11-13
	float scaleY, scaleZ;          // words 14-15
};

// 36 bytes, 4-byte aligned: the shading kernel reads one per hit and loads just the words it uses; the flags share one word
struct ShapeInfo {
	int32_t materialId, emitterId;
	uint32_t triOffset, vertOffset;
	uint32_t nTri, nVert;
	uint32_t cdfOffset;       // into DeviceScene::areaCdf (emitter meshes only); spheres: sphere-table index; cylinders / instances: DeviceScene::instances
	float invSurfaceArea;     // TriMesh::m_invSurfaceArea (trimesh.cpp:281)
	uint32_t hasNormals : 1;
	uint32_t hasTexcoords : 1;
	uint32_t hasTangents : 1; // per-vertex dpdu / dpdv present (TriMesh::computeTangentSpaceBasis ran for this mesh)
	uint32_t isSphere : 1;    // analytic sphere (src/shapes/sphere.cpp): one primitive, no mesh arrays
	uint32_t isCylinder : 1;  // analytic cylinder (src/shapes/cylinder.cpp): one primitive
	uint32_t isInstance : 1;  // instance of a shape group (src/shapes/instance.cpp): one primitive
};

// Instance of a shape group: both matrices of its Transform, and which group (DeviceScene::groups) it shows
struct InstanceDev { float w2o[16]; float o2w[16]; uint32_t group; float radius, length; uint32_t pad; };   // radius / length: analytic cylinder

struct MaterialDev { uint32_t type; float p[16]; uint32_t tex[3]; };   // tex: 0 constant, k: textures[k-1]; slot 2 = mask opacity (p[13..15])
struct TextureDev { uint32_t type; float p[12]; uint32_t pad[3]; };
// p[]: POINT pos | SPOT pos, axis row, cutoffAngle, invTransitionWidth, cosBeamWidth, cosCutoffAngle |
//      DIRECTIONAL dir, diskRadius | CONSTANT sphere centre, radius   (derived values filled by rebuildLuminaireCdf)
struct EmitterDev { float intensity[3]; uint32_t shape; float pdf; float weight; uint32_t type; float p[13]; };

struct CameraDev {
	float rasterToCamera[16];
	float cameraToWorld[16];
	float nearClip, farClip, apertureRadius, focusDepth;
	float shutterOpen, shutterOpenTime;
	uint32_t orthographic, pad;       // src/cameras/orthographic.cpp instead of perspective.cpp
};

struct FilmDev {
	int width, height, cropX, cropY, cropW, cropH;
	float filterSizeX, filterSizeY;   // TabulatedFilter::m_size
	float factorX, factorY;           // TabulatedFilter::m_factor = 15 / size
	int border;                       // ceil(max(size) - 0.5), renderproc.cpp:142-144
};

struct DeviceScene {
	const float4 *triAccel;     // sorted leaf order, 3 x float4 per triangle
	const float4 *nodes;        // BVH4QNode as 4 x float4 (two 256-bit loads)
	const uint4 *triIdx;        // kd primitive order: {v0, v1, v2 (global vertex ids), shape}
	const float4 *positions;    // global vertex arrays, xyz + pad
	const float4 *normals;
	const float4 *shadeTris;    // per sorted triangle, 6 x float4 (96 B = three 32-byte sectors): {p0, shape} {p1, prim}
	                            // {p2, v0} {n0, v1} {n1, v2} {n2, -}: everything fillIntersection needs behind ONE index
	const uint32_t *sortedOf;   // kd primitive order (ShapeInfo::triOffset + prim) -> sorted triangle index
	const float4 *tangents;     // global vertex array of {dpdu, dpdv} pairs; only read for shapes with hasTangents
	const float2 *texcoords;    // global vertex array; only read for shapes with hasTexcoords
	const TextureDev *textures;
	const float4 *texels;       // mip pyramids of the bitmap textures (rgb + pad per texel)
	const uint4 *mipLevels;     // {offset into texels, width, height, -} per level; TextureDev::pad = {first level, count}
	const float *mipLut;        // MIPMap's EWA weight table, 128 entries
	const InstanceDev *instances;            // instances of shape groups (empty unless hasInstances)
	const struct DeviceScene *groups;        // the groups' own scenes: acceleration structure and shading records in object space
	uint32_t hasInstances;
	uint32_t oneBits;           // 0x3F800000: the node decode's PRMT takes it from the parameter bank (trace.cuh: qPlane)
	uint32_t *traceFlags;       // bit 0: a traversal stack overflowed (the call that sees it fails with MTSGPU_ERR_LIMIT); bit 1: sampling.cuh
	int stackFast, stackTotal;  // traversal stack entries: shared-memory part, whole (TRACE_SMEM_STACK, + TRACE_LOCAL_STACK; tests shrink them)
	const float *emitterXf;     // per emitter: luminaireToWorld 4x4, worldToLuminaire 4x4 (collimated beams)
	const float *envData;       // environment map: worldToLuminaire 3x3, luminaireToWorld 3x3, density resolution, CDF, pdf (sampling.cuh)
	const struct CameraDev *camera;   // the render's camera: ray differentials of a camera-ray hit are rebuilt from it
	const ShapeInfo *shapes;
	const MaterialDev *materials;
	const EmitterDev *emitters;
	const float *lumCdf;        // nEmitters + 1 (DiscretePDF::m_cdf, pdf.h:82-100)
	const float *areaCdf;       // per emitter mesh: nTri + 1 entries at ShapeInfo::cdfOffset
	uint32_t nEmitters, nShapes, nTris, rootIsLeaf;
	uint32_t hasSpheres;        // some primitives are analytic spheres: traversal kernels take the variant that tests them
	int32_t background;         // index of the background luminaire (Scene::getBackgroundLuminaire), -1: none
	float lumOriginalSum;       // DiscretePDF::getOriginalSum()
	float aabbMin[3], aabbMax[3]; // enlarged kd-tree box (gkdtree.h:1172-1176)
};

// ---- transcendental functions -------------------------------------------------------------------
// The reference calls glibc's float routines (sinf, cosf, expf, logf, powf: correctly rounded in nearly
// all cases since glibc 2.28). CUDA's float versions are 1-2 ulp off, and a one-ulp difference in a
// sampled direction is amplified ~100x by a sharp Beckmann lobe. Evaluating in fp64 and rounding once
// reproduces the correctly rounded fp32 result, so sampled directions match the reference bit for bit
// in almost every case. These run a handful of times per shading event, never in traversal.
__device__ __forceinline__ float mts_sin(float x) { return (float) sin((double) x); }
__device__ __forceinline__ float mts_cos(float x) { return (float) cos((double) x); }
__device__ __forceinline__ float mts_tan(float x) { return (float) tan((double) x); }
__device__ __forceinline__ float mts_atan(float x) { return (float) atan((double) x); }
__device__ __forceinline__ float mts_atan2(float y, float x) { return (float) atan2((double) y, (double) x); }
__device__ __forceinline__ float mts_acos(float x) { return (float) acos((double) x); }
__device__ __forceinline__ float mts_log(float x) { return (float) log((double) x); }
__device__ __forceinline__ float mts_exp(float x) { return (float) exp((double) x); }
__device__ __forceinline__ float mts_pow(float x, float y) { return (float) pow((double) x, (double) y); }
// powf(x, 4.0f) (the Beckmann normalisation, roughmetal.cpp:77-78): x*x is exact in fp64 and its square is rounded once, so the
// result is the correctly rounded x^4 — what the general fp64 pow() gives, for 3 instructions instead of ~160 (ncu on k_shade:
// the pow() body alone was 8.3 % of the kernel's issue slots on the C4 scene)
__device__ __forceinline__ double mts_pow4d(float x) { const double d = (double) x * (double) x; return d * d; }
__device__ __forceinline__ float mts_pow4(float x) { return (float) mts_pow4d(x); }
// sine and cosine of one angle share the argument reduction
__device__ __forceinline__ void mts_sincos(float x, float &s, float &c) { double ds, dc; sincos((double) x, &ds, &dc); s = (float) ds; c = (float) dc; }

// ---- small vector helpers (shading math; not bit-exactness critical) ----------------------
struct V3 { float x, y, z; };
__host__ __device__ inline V3 v3(float x, float y, float z) { V3 r = { x, y, z }; return r; }
__host__ __device__ inline V3 v3(float4 a) { V3 r = { a.x, a.y, a.z }; return r; }
__host__ __device__ inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
__host__ __device__ inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ inline V3 cross(V3 a, V3 b) {
	return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__host__ __device__ inline float length(V3 a) { return sqrtf(dot(a, a)); }
// TVector3::operator/ multiplies by the reciprocal (core/vector.h:311-318)
__host__ __device__ inline V3 normalize(V3 a) { float r = 1.0f / length(a); return a * r; }
__host__ __device__ inline float maxc(V3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
__host__ __device__ inline bool isZero(V3 a) { return a.x == 0 && a.y == 0 && a.z == 0; }

// Frame(n) built by coordinateSystem (src/libcore/util.cpp:602-611, core/frame.h:35-83)
struct Frame3 { V3 s, t, n; };
__host__ __device__ inline Frame3 makeFrame(V3 n) {
	Frame3 f; f.n = n;
	if (fabsf(n.x) > fabsf(n.y)) {
		float invLen = 1.0f / sqrtf(n.x * n.x + n.z * n.z);
		f.s = v3(-n.z * invLen, 0.0f, n.x * invLen);
	} else {
		float invLen = 1.0f / sqrtf(n.y * n.y + n.z * n.z);
		f.s = v3(0.0f, -n.z * invLen, n.y * invLen);
	}
	f.t = cross(n, f.s);
	return f;
}
__host__ __device__ inline V3 toLocal(const Frame3 &f, V3 v) { return v3(dot(v, f.s), dot(v, f.t), dot(v, f.n)); }
__host__ __device__ inline V3 toWorld(const Frame3 &f, V3 v) { return f.s * v.x + f.t * v.y + f.n * v.z; }
