/*
 * mtsgpu.h — C ABI of libmtscuda.so, the CUDA (sm_100a) implementation of Mitsuba 0.2.1's
 * Monte Carlo path-tracing hot path. This is the drop-in boundary: the only things that
 * cross it are plain pointers, sizes and POD structs. The caller is the `gpupath`
 * SampleIntegrator plugin (mitsuba-renderer_b200/plugin/gpupath.cpp) or any FFI binding
 * (ctypes binding: mitsuba-renderer_b200/api.py).
 *
 * Citations are file:line under the reference tree (tomka/mitsuba-renderer).
 *
 * Conventions
 *   - every function returns 0 on success, a negative mtsgpu_status on failure;
 *     mtsgpu_last_error(ctx) returns a human-readable message for the last failure
 *   - the caller owns every host buffer; the library copies during the call
 *   - pointers named d_* are DEVICE pointers on the context's GPU; all others are host
 *   - matrices are row-major float[16] acting on column vectors (core/transform.h:133-149)
 *   - one host thread drives one context; `cancel` is the only cross-thread input
 *   - there is no CPU fallback: without a usable CUDA device every call fails
 */
#ifndef MTSGPU_H
#define MTSGPU_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MTSGPU_API __attribute__((visibility("default")))

typedef struct mtsgpu_ctx mtsgpu_ctx;

enum mtsgpu_status {
	MTSGPU_OK = 0,
	MTSGPU_ERR_CUDA = -1,       /* a CUDA runtime call failed */
	MTSGPU_ERR_ARG = -2,        /* invalid argument / call order */
	MTSGPU_ERR_UNSUPPORTED = -3,/* scene feature outside the hot path (non-mesh shape, texture, medium ...) */
	MTSGPU_ERR_CANCELLED = -4,  /* *cancel became non-zero during mtsgpu_render */
	MTSGPU_ERR_LIMIT = -5       /* a documented size limit was exceeded */
};

/* BSDF models (SURVEY.md "things a fresh reader must know" #3 maps them to north_star names) */
enum mtsgpu_bsdf_type {
	MTSGPU_BSDF_LAMBERTIAN = 0, /* src/bsdfs/lambertian.cpp:66   p[0..2] reflectance */
	MTSGPU_BSDF_MICROFACET = 1, /* src/bsdfs/microfacet.cpp:35   p[0..2] diffuseReflectance, p[3..5] specularReflectance,
	                               p[6] diffuseAmount, p[7] specularAmount, p[8] alphaB, p[9] intIOR, p[10] extIOR */
	MTSGPU_BSDF_MIRROR = 2,     /* src/bsdfs/mirror.cpp:26       p[0..2] specularReflectance */
	MTSGPU_BSDF_ROUGHMETAL = 3, /* src/bsdfs/roughmetal.cpp:33   p[0..2] specularReflectance, p[3] alphaB, p[4..6] ior, p[7..9] k */
	MTSGPU_BSDF_DIELECTRIC = 4, /* src/bsdfs/dielectric.cpp:47   p[0..2] specularReflectance, p[3..5] specularTransmittance,
	                               p[6] intIOR, p[7] extIOR */
	MTSGPU_BSDF_PHONG = 5,      /* src/bsdfs/phong.cpp:30        p[0..2] diffuseReflectance, p[3..5] specularReflectance, then the
	                               values AFTER Phong::configure (phong.cpp:79-99), as Phong::serialize writes them:
	                               p[6] diffuseAmount, p[7] specularAmount, p[8] exponent, p[9] specularSamplingWeight,
	                               p[10] diffuseSamplingWeight */
	MTSGPU_BSDF_ROUGHGLASS = 6, /* src/bsdfs/roughglass.cpp:81   p[0..2] specularReflectance, p[3..5] specularTransmittance,
	                               p[6] value of the constant alpha texture as RoughGlass::serialize writes it (for the
	                               phong distribution that is the exponent 2/alpha^2 - 2, roughglass.cpp:132-138),
	                               p[7] intIOR, p[8] extIOR, p[9] distribution: 0 beckmann, 1 phong, 2 ggx
	                               (roughglass.cpp:84-91); p[10] is ignored by the library (scene files keep the
	                               user-facing roughness of a phong distribution there for the oracle harness) */
	MTSGPU_BSDF_DIFFTRANS = 8,  /* src/bsdfs/difftrans.cpp:62    p[0..2] transmittance: ideally diffuse transmitter, two-sided */
	MTSGPU_BSDF_WARD = 7,       /* src/bsdfs/ward.cpp:42         p[0..2] diffuseReflectance, p[3..5] specularReflectance, then the
	                               values after Ward::configure (ward.cpp:120-139) as Ward::serialize writes them: p[6]
	                               diffuseAmount, p[7] specularAmount, p[8] alphaX, p[9] alphaY, p[10]
	                               specularSamplingWeight, p[11] diffuseSamplingWeight, p[12] model: 0 ward, 1 ward-duer,
	                               2 balanced (ward.cpp:45-52). alphaX != alphaY makes the BSDF anisotropic: its meshes must
	                               carry texture coordinates and vertex normals, and the library computes their tangents as
	                               TriMesh::computeTangentSpaceBasis does (src/librender/trimesh.cpp:547-662) */
	MTSGPU_BSDF_TYPE_LAST = 8
};

/* OR-ed into mtsgpu_material.type: the BRDF is wrapped in the `twosided` adapter (src/bsdfs/twosided.cpp:36-143),
 * i.e. queries with wi.z < 0 are mirrored to the upper side. Not valid for dielectric (the adapter rejects
 * BSDFs with transmission, twosided.cpp:72-74). */
#define MTSGPU_BSDF_TWOSIDED  0x100u
/* OR-ed into mtsgpu_material.type: the BSDF (with its twosided adapter, if any) is nested in a `mask` BSDF
 * (src/bsdfs/mask.cpp:28-160): f and pdf are scaled by the luminance of the opacity spectrum in p[13..15] (texture slot 2),
 * and sampling picks, with probability 1 - opacity, a delta transmission straight through the surface. */
#define MTSGPU_BSDF_MASKED    0x200u
#define MTSGPU_BSDF_TYPE_MASK 0x0FFu

typedef struct mtsgpu_material {
	uint32_t type;   /* mtsgpu_bsdf_type, optionally | MTSGPU_BSDF_TWOSIDED | MTSGPU_BSDF_MASKED */
	float p[16];
	uint32_t texture[3]; /* 0: the spectrum in p[0..2] (slot 0) / p[3..5] (slot 1) / p[13..15] (slot 2: the opacity of a MASK
	                        wrapper) is constant; k > 0: it is read from
	                        texture k-1 of mtsgpu_upload_textures at the hit's uv (Texture::getValue(its), e.g.
	                        lambertian.cpp:96-102). Slots that the reference keeps as plain spectra (mirror, dielectric,
	                        roughmetal eta/k) cannot be textured. */
} mtsgpu_material;

/* 2D textures (src/librender/texture.cpp:43-82: uv' = uv * uvScale + uvOffset, then the texture's own lookup).
 *   CHECKERBOARD  src/textures/checkerboard.cpp:28-55   p[0..2] brightColor, p[3..5] darkColor, p[6..7] uvOffset,
 *                                                       p[8..9] uvScale
 *   GRID          src/textures/gridtexture.cpp:29-60    p[0..2] brightColor, p[3..5] darkColor (the lines), p[6..7] uvOffset,
 *                                                       p[8..9] uvScale, p[10] lineWidth
 *   BITMAP        src/textures/ldrtexture.cpp:35-253 (and exrtexture.cpp) over MIPMap, src/librender/mipmap.cpp:
 *                 p[0] filterType (0 ewa, 1 trilinear, 2 none: MIPMap::EFilterType, mipmap.h:39-46), p[1] wrapMode
 *                 (0 repeat, 1 black, 2 white, 3 clamp: mipmap.h:32-37), p[2] maxAnisotropy, p[6..7] uvOffset,
 *                 p[8..9] uvScale. The pixels follow through mtsgpu_upload_bitmap. At the hit of a camera ray the
 *                 lookup is filtered with the uv partials of the ray differentials (Intersection::computePartials,
 *                 src/librender/intersection.cpp:5-75); every later hit takes the bilinear level-0 lookup, which is
 *                 what the reference's filters reduce to for a ray without differentials. */
enum mtsgpu_texture_type { MTSGPU_TEXTURE_CHECKERBOARD = 0, MTSGPU_TEXTURE_BITMAP = 1, MTSGPU_TEXTURE_GRID = 2 };
enum mtsgpu_bitmap_format { MTSGPU_BITMAP_U8 = 0, MTSGPU_BITMAP_F32 = 1 };
typedef struct mtsgpu_texture {
	uint32_t type;   /* mtsgpu_texture_type */
	float p[12];
} mtsgpu_texture;

/* Luminaires. List them in the order of Scene::getLuminaires(), which is the order of the luminaire CDF
 * (src/librender/scene.cpp:320-330): luminaires that are scene children come first, in the order they were added
 * (scene.cpp:604-612); area luminaires follow, in the order Scene::initialize meets their shapes
 * (scene.cpp:296-299, 656-664). Area luminaires must therefore appear in the order of their shapes.
 *   AREA         src/luminaires/area.cpp:35-83         bound to the mesh whose emitterId names it; p unused
 *   POINT        src/luminaires/point.cpp:27-66        p[0..2] position
 *   SPOT         src/luminaires/spot.cpp:32-127        p[0..2] position, p[3..5] third row of the 3x3 part of
 *                                                      worldToLuminaire (cosTheta = p[3..5] . d), p[6] beamWidth and
 *                                                      p[7] cutoffAngle in radians (as SpotLuminaire::serialize writes
 *                                                      them); constant projection texture only
 *   DIRECTIONAL  src/luminaires/directional.cpp:29-90  p[0..2] direction (normalised), p[3] diskRadius; 0 = derive it
 *                                                      from the scene's bounding sphere as preprocess() does
 *   CONSTANT     src/luminaires/constant.cpp:27-86     background; p[0..2] centre and p[3] radius of its bounding
 *                                                      sphere; radius 0 = derive it as preprocess() does (scene
 *                                                      sphere x 1.01, grown to hold the camera). At most one.
 *   ENVMAP       src/luminaires/envmap.cpp:38-199      background from a latitude-longitude map, importance-sampled
 *                                                      through a coarse MIP level; p[0..3] bounding sphere as for
 *                                                      CONSTANT, p[4] intensityScale; intensity unused. The map and
 *                                                      its transform follow through mtsgpu_upload_envmap. CONSTANT and
 *                                                      ENVMAP exclude each other (one background luminaire).
 *   COLLIMATED   src/luminaires/collimated.cpp:27-75   a beam of parallel light: p[0] radius; its toWorld transform follows
 *                                                      through mtsgpu_upload_emitter_transform. Lights the points whose
 *                                                      projection on the beam's plane lies inside the disc, in front of it. */
enum mtsgpu_emitter_type {
	MTSGPU_EMITTER_AREA = 0, MTSGPU_EMITTER_POINT = 1, MTSGPU_EMITTER_SPOT = 2, MTSGPU_EMITTER_DIRECTIONAL = 3,
	MTSGPU_EMITTER_CONSTANT = 4, MTSGPU_EMITTER_ENVMAP = 5, MTSGPU_EMITTER_COLLIMATED = 6
};
typedef struct mtsgpu_emitter {
	float intensity[3];
	float samplingWeight; /* Luminaire::getSamplingWeight(), luminaire.h:180; weight in the luminaire CDF */
	uint32_t type;        /* mtsgpu_emitter_type */
	float p[11];
} mtsgpu_emitter;

enum mtsgpu_mode { MTSGPU_MODE_PATH = 0, MTSGPU_MODE_DIRECT = 1 };

/* Integrator + sampler properties (src/librender/integrator.cpp:270-295, src/integrators/direct/direct.cpp:34-36,
 * src/samplers/independent.cpp:34) and the sample range this context renders. */
typedef struct mtsgpu_render_params {
	int32_t mode;             /* mtsgpu_mode: MIPathTracer::Li (path.cpp:47-216) or MIDirectIntegrator::Li (direct.cpp:63-203) */
	int32_t maxDepth;         /* -1 = unbounded */
	int32_t rrDepth;          /* Russian roulette from this depth (default 10) */
	int32_t strictNormals;
	int32_t luminaireSamples; /* direct only */
	int32_t bsdfSamples;      /* direct only */
	uint32_t sppTotal;        /* sampler sampleCount for the whole image */
	uint32_t sppBegin;        /* this call renders samples [sppBegin, sppEnd) of every pixel: */
	uint32_t sppEnd;          /*   the multi-GPU shard; 0..sppTotal on one GPU */
	uint32_t seed;            /* counter-based RNG key; streams are keyed by (pixel, global sample index) */
	uint32_t importanceSampleLuminaires; /* Scene property, scene.cpp:51; default 1 */
	uint32_t samplerType;     /* mtsgpu_sampler_type */
	uint32_t samplerDepth;    /* ldsampler `depth` property (ldsampler.cpp:59): the first samplerDepth 1D and 2D draws of
	                             every pixel sample come from the low-discrepancy sets, later ones are independent */
	uint32_t reserved;
} mtsgpu_render_params;

/* MTSGPU_SAMPLER_INDEPENDENT: src/samplers/independent.cpp:72-81.
 * MTSGPU_SAMPLER_LD: src/samplers/ldsampler.cpp:104-196 — per pixel and dimension a (0,2)-sequence (van der Corput /
 * Sobol' radical inverses, XOR-scrambled) visited in a random order; sppTotal must be a power of two (the reference
 * rounds the sample count up, ldsampler.cpp:61-65). Scrambles and orders are counter-based here too (keyed by pixel,
 * dimension, seed), so the image does not depend on batching or on the GPU count. */
/* MTSGPU_SAMPLER_STRATIFIED: src/samplers/stratified.cpp:123-182 — per pixel and draw the R x R strata of the unit square
 * (R*R = sppTotal, which must be a perfect square; the reference rounds up, stratified.cpp:37-47) are visited in a
 * random order with a jittered point in each; 1D draws use R*R strata of the unit interval. */
/* MTSGPU_SAMPLER_HAMMERSLEY: src/samplers/hammersley.cpp:72-99 — as HALTON, except that the first value of pixel sample i
 * is i / sampleCount and value d > 0 is radicalInverse(prime d-1, i). */
/* MTSGPU_SAMPLER_HALTON: src/samplers/halton.cpp:68-84 — value number d of pixel sample i is radicalInverse(prime d, i)
 * (src/libcore/util.cpp:740-750), the same for every pixel; fully deterministic, so images can be compared with the
 * reference pixel by pixel. Like the reference it supports no sample arrays (luminaireSamples = bsdfSamples = 1). */
enum mtsgpu_sampler_type { MTSGPU_SAMPLER_INDEPENDENT = 0, MTSGPU_SAMPLER_LD = 1, MTSGPU_SAMPLER_STRATIFIED = 2,
                           MTSGPU_SAMPLER_HALTON = 3, MTSGPU_SAMPLER_HAMMERSLEY = 4 };

typedef struct mtsgpu_stats {
	uint64_t cameraRays;      /* "Camera ray generations" (perspective.cpp:24) */
	uint64_t normalRays;      /* closest-hit queries, "Normal rays traced" (skdtree.cpp:40) */
	uint64_t shadowRays;      /* any-hit queries,   "Shadow rays traced" (skdtree.cpp:41) */
	uint64_t samples;         /* completed pixel samples */
	uint64_t kernelLaunches;  /* kernels launched by this call */
	double   traceClosestMs;  /* device time spent in the closest-hit kernel (CUDA events) */
	double   traceAnyMs;      /* ... in the any-hit kernel */
	double   shadeMs;         /* ... in ray generation + shading + splatting */
	double   totalMs;         /* whole call, device time */
	/* filled only when counters are enabled (mtsgpu_set_counters): traversal work on the render's own ray queues, the
	 * counterpart of rayIntersectHavranCollectStatistics (include/mitsuba/render/sahkdtree3.h:310-410) */
	uint64_t closestNodeVisits, closestTriTests, shadowNodeVisits, shadowTriTests;
} mtsgpu_stats;

/* Progress and preview while mtsgpu_render / mtsgpu_render_device run. Replaces, for this integrator, the ProgressReporter
 * updates and RenderQueue::signalRefresh / the mutex-guarded film merge that BlockedRenderProcess::processResult performs after
 * every block (src/librender/renderproc.cpp:123-130, include/mitsuba/render/renderqueue.h:109-112): between wavefront batches,
 * at most once per `minIntervalSeconds`, `fn(user, fraction, film)` is called ON THE THREAD THAT CALLED mtsgpu_render with the
 * fraction of the work enqueued so far and — if wantFilm — the {sum spec.rgb, sum alpha, sum weight} accumulators as they stood
 * at that point (cropW * cropH * 5 floats in pinned host memory owned by the library, valid during the call; tiles not yet
 * rendered are zero). The copy runs on its own stream from a device-side snapshot, the render is not synchronised for it.
 * A last call with fraction 1 and film = NULL follows the render. fn = NULL switches it off. */
typedef void (*mtsgpu_progress_fn)(void *user, double fraction, const float *film);
MTSGPU_API int mtsgpu_set_progress(mtsgpu_ctx *ctx, mtsgpu_progress_fn fn, void *user, double minIntervalSeconds, int wantFilm);

/* Closest-hit record returned by mtsgpu_trace; mirrors the reference's IntersectionCache
 * (include/mitsuba/render/skdtree.h:232-236) plus t. Miss: t = +inf, shape = prim = 0xFFFFFFFF. */
typedef struct mtsgpu_hit {
	float t, u, v;
	uint32_t shape, prim;
} mtsgpu_hit;

enum mtsgpu_trace_mode {
	MTSGPU_TRACE_CLOSEST = 0, /* ShapeKDTree::rayIntersect(const Ray&, Intersection&), skdtree.cpp:108-132 */
	MTSGPU_TRACE_ANY = 1      /* ShapeKDTree::rayIntersect(const Ray&), skdtree.cpp:180-199 (shadow rays) */
};

/* ---- life cycle ---------------------------------------------------------------------------- */

/* Create a context on CUDA device `device`. Replaces nothing in the reference (it has no device). */
MTSGPU_API int mtsgpu_create(int device, mtsgpu_ctx **out);
MTSGPU_API void mtsgpu_destroy(mtsgpu_ctx *ctx);
MTSGPU_API const char *mtsgpu_last_error(const mtsgpu_ctx *ctx);
MTSGPU_API const char *mtsgpu_version(void);

/* ---- scene upload: what the plugin reads through Scene's public getters ------ */

/* BSDF parameter blocks, decoded by the plugin from BSDF::serialize (e.g. roughmetal.cpp:169-176). */
MTSGPU_API int mtsgpu_upload_materials(mtsgpu_ctx *ctx, const mtsgpu_material *materials, int count);

/* Scene::getLuminaires() (include/mitsuba/render/scene.h:517-519) restricted to AreaLuminaire
 * (src/luminaires/area.cpp:35-83): intensity + Luminaire::getSamplingWeight. */
/* Textures referenced by mtsgpu_material.texture (Texture2D::getValue, src/librender/texture.cpp:73-82). */
MTSGPU_API int mtsgpu_upload_textures(mtsgpu_ctx *ctx, const mtsgpu_texture *textures, int count);
/* Pixels of BITMAP texture `textureIndex` (after mtsgpu_upload_textures, before mtsgpu_build_accel), top row first.
 * MTSGPU_BITMAP_U8: `channels` (1, 2, 3 or 4; one and two take the first as grey) bytes per pixel as Bitmap::getData() holds them; they are linearised with
 * the table of LDRTexture::initializeFrom (src/textures/ldrtexture.cpp:124-207): gamma == -1 selects the sRGB curve,
 * any other value is the exponent. MTSGPU_BITMAP_F32: `channels` (3 or 4) linear floats per pixel, as EXRTexture hands
 * them to MIPMap::fromBitmap (src/textures/exrtexture.cpp:52-60); gamma is ignored. The library builds the pyramid of
 * MIPMap::MIPMap (src/librender/mipmap.cpp:31-137: Lanczos up-sampling to powers of two, box-filtered levels). */
MTSGPU_API int mtsgpu_upload_bitmap(mtsgpu_ctx *ctx, int textureIndex, int width, int height, int channels,
	const void *pixels, int format, float gamma);
/* The map of ENVMAP luminaire `emitterIndex` (after mtsgpu_upload_emitters, before mtsgpu_build_accel): `channels` (3 or
 * 4) linear floats per pixel, top row first, as Bitmap::loadEXR hands them to MIPMap::fromBitmap
 * (src/luminaires/envmap.cpp:44-47; EWA pyramid, `repeat` wrapping), and the luminaire's toWorld transform with its
 * inverse (Luminaire::m_luminaireToWorld / m_worldToLuminaire; only the 3x3 parts are used, on directions). */
MTSGPU_API int mtsgpu_upload_envmap(mtsgpu_ctx *ctx, int emitterIndex, int width, int height, int channels,
	const float *pixels, const float luminaireToWorld[16], const float worldToLuminaire[16]);
/* luminaireToWorld and its inverse (Luminaire::m_luminaireToWorld / m_worldToLuminaire, src/librender/luminaire.cpp:28-34)
 * of luminaire `emitterIndex`, for the types that need a full transform (COLLIMATED). After mtsgpu_upload_emitters. */
MTSGPU_API int mtsgpu_upload_emitter_transform(mtsgpu_ctx *ctx, int emitterIndex, const float luminaireToWorld[16],
	const float worldToLuminaire[16]);
MTSGPU_API int mtsgpu_upload_emitters(mtsgpu_ctx *ctx, const mtsgpu_emitter *emitters, int count);

/* One TriMesh (include/mitsuba/render/trimesh.h:100-133) at kd-tree shape index `shapeIndex`
 * (ShapeKDTree::addShape order, skdtree.cpp:45-60). Shapes must be uploaded in index order 0,1,2...
 * positions: 3*nVert floats; normals: 3*nVert floats or NULL (=> shading frame = geometric frame,
 * skdtree.h:405-408); texcoords: 2*nVert floats or NULL (its.uv = 0, skdtree.h:414).
 * materialId < 0: shape has no BSDF => not an occluder (shape.h:361) and paths end on it (path.cpp:70-77).
 * emitterId  < 0: not a luminaire. */
MTSGPU_API int mtsgpu_upload_mesh(mtsgpu_ctx *ctx, int shapeIndex,
	const float *positions, const float *normals, const float *texcoords, uint32_t nVert,
	const uint32_t *triangles, uint32_t nTri, int materialId, int emitterId);

/* PerspectiveCameraImpl (src/cameras/perspective.cpp:43-112): m_rasterToCamera, m_cameraToWorld,
 * clip planes, thin-lens parameters, shutter. */
/* Analytic sphere as shape `shapeIndex` (src/shapes/sphere.cpp:31-173): the fields Sphere::serialize writes
 * (sphere.cpp:71-77) — objectToWorld with the scale removed and its inverse (row-major 4x4), radius, centre,
 * `inverted`. It is one primitive of the acceleration structure; hits report prim = 0xFFFFFFFF (KNoTriangleFlag,
 * include/mitsuba/render/triaccel.h:27) and u = v = 0. With emitterId >= 0 the sphere carries that area
 * luminaire: it is then sampled by solid angle — the cone it subtends from outside, uniformly by area from inside
 * (Sphere::sampleSolidAngle / pdfSolidAngle, src/shapes/sphere.cpp:191-247, through AreaLuminaire::sample,
 * src/luminaires/area.cpp:71-86). */
MTSGPU_API int mtsgpu_upload_sphere(mtsgpu_ctx *ctx, int shapeIndex, const float objectToWorld[16], const float worldToObject[16],
                                    float radius, const float center[3], int inverted, int materialId, int emitterId);

/* An analytic cylinder (src/shapes/cylinder.cpp:30-205) at kd-tree shape index `shapeIndex`: the open tube
 * x^2 + y^2 = radius^2, 0 <= z <= length of its object space (no caps), under objectToWorld (scale already taken out,
 * cylinder.cpp:52-58; worldToObject is its inverse). One primitive, tested as Cylinder::rayIntersect does (ray to
 * object space, quadratic in x / y, the nearer root inside the z range wins); hits report prim = 0xFFFFFFFF, u = v = 0.
 * Cannot carry an area luminaire: the reference's own Cylinder has no pdfArea ("Cylinder::pdfArea(): Not implemented!",
 * src/librender/shape.cpp via Shape::pdfSolidAngle), so a render that reaches such a luminaire stops with an error there. */
MTSGPU_API int mtsgpu_upload_cylinder(mtsgpu_ctx *ctx, int shapeIndex, const float objectToWorld[16], const float worldToObject[16],
	float radius, float length, int materialId);

/* Instancing (src/shapes/shapegroup.cpp, src/shapes/instance.cpp). A shape group is a set of triangle meshes with an
 * acceleration structure of its own (ShapeGroup::m_kdtree) that is not part of the scene by itself; mesh `meshIndex`
 * of group `groupId` (both 0, 1, 2 ... in order) is uploaded like a scene mesh, in the group's object space. Luminaires
 * cannot be instanced ("Instancing of luminaires is not supported", shapegroup.cpp:66-67). */
MTSGPU_API int mtsgpu_upload_group_mesh(mtsgpu_ctx *ctx, int groupId, int meshIndex, const float *positions, const float *normals,
	const float *texcoords, uint32_t nVert, const uint32_t *triangles, uint32_t nTri, int materialId);
/* An <shape type="instance"> at kd-tree shape index `shapeIndex`: group `groupId` under objectToWorld (worldToObject is
 * its inverse as the reference's Transform holds it). One primitive of the scene's acceleration structure, boxed like
 * Instance::getAABB (instance.cpp:47-56); a ray that reaches it is taken to object space — direction not renormalised,
 * so t is shared (Instance::rayIntersect, instance.cpp:72-86) — and continues in the group's own structure, inside
 * the interval clipped to the group's box (ShapeKDTree::rayIntersect, skdtree.h:435-466). Hits report the instance's
 * shape index and prim = 0xFFFFFFFF; shading follows Instance::fillIntersectionRecord (instance.cpp:88-98). */
MTSGPU_API int mtsgpu_upload_instance(mtsgpu_ctx *ctx, int shapeIndex, int groupId, const float objectToWorld[16],
	const float worldToObject[16]);

MTSGPU_API int mtsgpu_set_camera(mtsgpu_ctx *ctx, const float rasterToCamera[16], const float cameraToWorld[16],
	float nearClip, float farClip, float apertureRadius, float focusDepth, float shutterOpen, float shutterOpenTime);

/* OrthographicCamera (src/cameras/orthographic.cpp:30-119) instead: rasterToCamera = cameraToScreen.inverse() *
 * rasterToScreen with cameraToScreen = Transform::orthographic(nearClip, farClip) (orthographic.cpp:83-86); rays start at
 * cameraToWorld(rasterToCamera(x, y, 0)) and run along cameraToWorld(0, 0, 1) for t in [0, farClip - nearClip]
 * (orthographic.cpp:104-119). No lens sample (needsLensSample() is false).
 * Parity note: in the reference this class re-declares m_nearClip / m_farClip (orthographic.cpp:176), and it is those
 * uninitialised copies that its constructor and generateRay read: as compiled here every ray comes out with
 * maxt = 0 and the reference renders nothing. Ray origins and directions agree with it (tested); the interval is the
 * intended one. */
MTSGPU_API int mtsgpu_set_camera_orthographic(mtsgpu_ctx *ctx, const float rasterToCamera[16], const float cameraToWorld[16],
	float nearClip, float farClip, float shutterOpen, float shutterOpenTime);

/* Film geometry (src/librender/film.cpp:24-51) and the film's TabulatedFilter
 * (include/mitsuba/render/rfilter.h:65-118): the 16x16 table m_values, and the filter half-size. */
MTSGPU_API int mtsgpu_set_film(mtsgpu_ctx *ctx, int width, int height, int cropX, int cropY, int cropW, int cropH,
	const float filterTable[256], float filterSizeX, float filterSizeY);

/* Replaces ShapeKDTree::build (skdtree.cpp:62-106): TriAccel precompute (triaccel.h:63-96) and a
 * GPU-built wide BVH instead of the SAH kd-tree (gkdtree.h:913-1214). Also builds the luminaire and
 * per-mesh area CDFs (scene.cpp:320-330, trimesh.cpp:266-283). */
MTSGPU_API int mtsgpu_build_accel(mtsgpu_ctx *ctx);

/* ---- the hot path ---------------------------------------------------------------------------- */

/* Replaces SampleIntegrator::render -> BlockedRenderProcess -> renderBlock -> Li -> putSample ->
 * Film::putImageBlock (integrator.cpp:87-268, renderproc.cpp:106-160, imageblock.h:80-138).
 * outFilm: HOST buffer of cropW*cropH*5 floats, per pixel {sum spec R,G,B, sum alpha, sum weight}
 * (the un-normalised accumulators of exrfilm.cpp:37-44), overwritten.
 * cancel may be NULL; polled between wavefront iterations (Integrator::cancel, integrator.cpp:82-85). */
MTSGPU_API int mtsgpu_render(mtsgpu_ctx *ctx, const mtsgpu_render_params *params, volatile const int *cancel,
	float *outFilm, mtsgpu_stats *stats);

/* Same, but the film stays on the device: d_outFilm is a DEVICE buffer (cropW*cropH*5 floats) that the
 * caller then reduces across GPUs (one ncclReduce; replaces the mutex-guarded merge of
 * BlockedRenderProcess::processResult, renderproc.cpp:123-130). */
MTSGPU_API int mtsgpu_render_device(mtsgpu_ctx *ctx, const mtsgpu_render_params *params, volatile const int *cancel,
	float *d_outFilm, mtsgpu_stats *stats);

/* Several GPUs, one process (the `mitsuba` CLI case). ctxs[i] are contexts on DIFFERENT devices that hold the
 * same scene, camera and film. The sample range [sppBegin, sppEnd) of `params` is split evenly over them and
 * rendered concurrently (one host thread per GPU); the film accumulators are then summed on ctxs[0]'s GPU by a
 * single kernel that loads the peers' buffers over NVLink peer mappings, and copied to outFilm (host). Replaces the
 * distribution of work units to local/remote workers and the mutex-guarded merge of their results
 * (BlockedRenderProcess, src/librender/renderproc.cpp:106-160; RemoteWorker, src/libcore/sched_remote.cpp).
 * stats: counters summed, times = max over GPUs (+ the reduce). A progress callback set on ctxs[0] (mtsgpu_set_progress) is
 * called on the CALLING thread as in mtsgpu_render, with the fraction of ctxs[0]'s share enqueued and film = NULL (partial
 * films stay on their GPUs); the other contexts' callbacks are not used. */
MTSGPU_API int mtsgpu_render_multi(mtsgpu_ctx **ctxs, int nCtx, const mtsgpu_render_params *params,
	volatile const int *cancel, float *outFilm, mtsgpu_stats *stats);

/* Batch ray queries against the scene; rays are rows of 8 floats {o.xyz, mint, d.xyz, maxt}
 * (include/mitsuba/core/ray.h:34-40 without dRcp/time). mode CLOSEST writes n mtsgpu_hit records,
 * mode ANY writes n uint32 (1 = occluded). Host buffers; copies are part of the call. */
MTSGPU_API int mtsgpu_trace(mtsgpu_ctx *ctx, const float *rays, size_t n, int mode, void *out);

/* Device-resident variant of mtsgpu_trace (same reference entry points, skdtree.cpp:108-132,180-199):
 * d_rays / d_out are DEVICE pointers; runs on the context's stream and returns after the kernel has completed. elapsedMs (may be NULL) receives the kernel's device time. */
MTSGPU_API int mtsgpu_trace_device(mtsgpu_ctx *ctx, const float *d_rays, size_t n, int mode, void *d_out,
	float *elapsedMs);

/* ---- per-function test hooks (parity tests call the device functions through these) -------- */

/* BSDF::f / pdf / sample (bsdf.h:246-292; bsdf.cpp:37-48). queries: rows of 9 floats
 * {materialId, wi.xyz, wo.xyz, sample.x, sample.y}; out: rows of 16 floats
 * {f.rgb, pdf | sampled wo.xyz, sampled f.rgb, sampled pdf, sampledType | 2-arg sample value rgb, 0}. */
MTSGPU_API int mtsgpu_bsdf_eval(mtsgpu_ctx *ctx, const float *queries, size_t n, float *out);

/* Scene::sampleLuminaire without the visibility test + Scene::pdfLuminaire (scene.cpp:381-415).
 * queries: rows {p.xyz, sample.x, sample.y}; out rows of 16 floats
 * {sRec.p, sRec.n, d, pdf, value.rgb, luminaireIndex (-1 = failed), pdfLuminaire, 0}. */
MTSGPU_API int mtsgpu_emitter_sample(mtsgpu_ctx *ctx, const float *queries, size_t n, float *out);

/* PerspectiveCameraImpl::generateRay (perspective.cpp:77-112). samples: rows {x, y, lens.x, lens.y};
 * out rows of 8 floats {o.xyz, mint, d.xyz, maxt}. */
MTSGPU_API int mtsgpu_camera_rays(mtsgpu_ctx *ctx, const float *samples, size_t n, float *out);

/* ImageBlock::putSample + Film::putImageBlock (imageblock.h:80-138, exrfilm.cpp:130-156).
 * samples: rows {x, y, r, g, b, alpha}; outFilm as in mtsgpu_render. */
MTSGPU_API int mtsgpu_film_splat(mtsgpu_ctx *ctx, const float *samples, size_t n, float *outFilm);

/* No counterpart in the reference. The same splat as mtsgpu_film_splat (host samples in, host film out,
 * identical arithmetic) for callers that are not allowed to allocate GPU memory while they run: a progress
 * or preview refresh on a latency-critical thread while a render already holds most of the device's memory
 * (cudaMalloc may fail there, and it stalls the device), or a host that fixes the library's memory budget up
 * front. This call performs no cudaMalloc/cudaFree and uses no allocating library routine: the staged
 * samples, the filter table and the film accumulators live in one fixed block of device memory that is part
 * of the library image, MTSGPU_STATIC_SCRATCH_BYTES per GPU, held from library load until process exit.
 *   - limit: 1024 + 20*cropW*cropH + 24*n bytes (256-byte rounding between the parts) must fit in the block;
 *     otherwise the call fails with MTSGPU_ERR_LIMIT and a message that names both sizes.
 *   - the first call on a device also looks up the block's device address (once per device); make that
 *     first call before starting any stream capture.
 *   - all calls on a GPU share the block and it keeps the previous call's contents; the call clears the
 *     accumulators itself, but two calls that overlap on the same GPU (different contexts or threads)
 *     race on it — serialise them. */
#define MTSGPU_STATIC_SCRATCH_BYTES (4u << 20)
MTSGPU_API int mtsgpu_film_splat_static(mtsgpu_ctx *ctx, const float *samples, size_t n, float *outFilm);

/* ---- `.serialized` triangle-mesh files (host side; the data format in front of the path) ------
 * A mesh as the reference's `serialized` shape would hand it to the kd-tree: file read
 * (TriMesh::TriMesh(Stream *, int), src/librender/trimesh.cpp:156-237), object-to-world transform and winding
 * (SerializedMesh::SerializedMesh, src/shapes/serialized.cpp:30-91), vertex normals (TriMesh::computeNormals,
 * src/librender/trimesh.cpp:472-540). Arrays are malloc'ed by the library and released by
 * mtsgpu_serialized_free. These three calls need no GPU and no context; errors are reported through
 * mtsgpu_last_error(NULL). */
typedef struct mtsgpu_mesh_data {
	uint32_t nVert, nTri;
	uint32_t flags;        /* the file's flag word (trimesh.cpp:72-80): 1 normals, 2 texcoords, 8 colors, 0x10 face normals,
	                          0x1000 / 0x2000 single / double precision */
	float *positions;      /* 3 per vertex */
	float *normals;        /* 3 per vertex, NULL when the mesh is shaded with face normals */
	float *texcoords;      /* 2 per vertex or NULL */
	float *colors;         /* 3 per vertex or NULL */
	uint32_t *triangles;   /* 3 per triangle */
} mtsgpu_mesh_data;
/* Number of meshes in the file: the trailing table written by the converter (src/converter/converter.cpp:293-296). */
MTSGPU_API int mtsgpu_serialized_count(const char *path, int *count);
/* toWorld / toWorldInv: row-major 4x4 or NULL (identity; a missing inverse is computed); faceNormals and
 * flipNormals are the shape properties of the same names (src/librender/trimesh.cpp:62-63). */
MTSGPU_API int mtsgpu_serialized_load(const char *path, int shapeIndex, const float *toWorld, const float *toWorldInv,
                                      int faceNormals, int flipNormals, mtsgpu_mesh_data *out);
/* Releases the arrays of a loaded mesh (what TriMesh::~TriMesh does, src/librender/trimesh.cpp:239-252). */
MTSGPU_API void mtsgpu_serialized_free(mtsgpu_mesh_data *mesh);
/* Writer: TriMesh::serialize(Stream *) per mesh (src/librender/trimesh.cpp:748-790, single precision, zlib) followed
 * by the offset table; the reference's loader reads the result. */
MTSGPU_API int mtsgpu_serialized_save(const char *path, const mtsgpu_mesh_data *meshes, int count);
/* mtsgpu_serialized_load + mtsgpu_upload_mesh: file to GPU scene without a detour through the caller
 * (the `serialized` shape inside a scene, src/shapes/serialized.cpp:30-91). */
MTSGPU_API int mtsgpu_upload_serialized(mtsgpu_ctx *ctx, int shapeIndex, const char *path, int fileShapeIndex,
                                        const float *toWorld, const float *toWorldInv, int faceNormals, int flipNormals,
                                        int materialId, int emitterId);

/* ---- film development and image files (what follows the path) --------------------------------
 * Film accumulators are cropH x cropW x {sum spec.rgb, sum alpha, sum weight} as mtsgpu_render returns them. */
enum mtsgpu_film_format {
	MTSGPU_FILM_EXR = 0,  /* EXRFilm::develop (src/films/exrfilm.cpp:156-210) + Bitmap::saveEXR (src/libcore/bitmap.cpp:599-628):
	                         half RGBA scanline OpenEXR file, ZIP blocks; option = the film's `alpha` property; no banner */
	MTSGPU_FILM_M = 1     /* MFilm::develop (src/films/mfilm.cpp:174-224): MATLAB matrix text; option = `spectra` property */
};
/* EXRFilm::develop's per-pixel normalisation on the GPU (exrfilm.cpp:163-176): rgba = cropH x cropW x 4 floats. */
MTSGPU_API int mtsgpu_film_develop(mtsgpu_ctx *ctx, const float *film, int hasAlpha, float *rgba);
/* The same with device pointers (exrfilm.cpp:163-176), for a film that mtsgpu_render_device left on the GPU. */
MTSGPU_API int mtsgpu_film_develop_device(mtsgpu_ctx *ctx, const float *d_film, int hasAlpha, float *d_rgba);
/* Film::develop(destFile) for the two film types above (exrfilm.cpp:156-210, mfilm.cpp:174-224): develops on the
 * GPU (film may be a host or a device pointer) and writes the file. */
MTSGPU_API int mtsgpu_film_save(mtsgpu_ctx *ctx, const float *film, int filmOnDevice, int format, int option, const char *path);
/* File writers alone (host code, no GPU, no context; errors through mtsgpu_last_error(NULL)):
 * Bitmap::saveEXR (src/libcore/bitmap.cpp:599-628) for width x height x 4 floats -> half A,B,G,R channels, and the text
 * layout of MFilm::develop (src/films/mfilm.cpp:183-221) for width x height x channels (1 or 3) values. */
MTSGPU_API int mtsgpu_write_exr(const char *path, const float *rgba, int width, int height);
MTSGPU_API int mtsgpu_write_mfilm(const char *path, const float *values, int width, int height, int channels);

/* Test hook for the samplers (src/samplers/independent.cpp:72-81, ldsampler.cpp:104-196): the value pair that 2D draw
 * number `draw` (0 = first next2D() of a pixel sample) returns for every sample 0..sppTotal-1 of pixel `pixel`
 * (= y * width + x); out holds sppTotal x 2 floats. */
MTSGPU_API int mtsgpu_sampler_values(mtsgpu_ctx *ctx, const mtsgpu_render_params *params, uint32_t pixel, uint32_t draw, float *out);

/* Closest-hit query reporting the interpolated texture coordinates (fillIntersectionRecord, skdtree.h:408-415) and the
 * two texture slots of the hit's material evaluated there: out rows of 8 floats {uv, slot 0 rgb, slot 1 rgb}; zeros
 * on a miss or on a shape without a BSDF. */
MTSGPU_API int mtsgpu_intersect_uv(mtsgpu_ctx *ctx, const float *rays, size_t n, float *out);

/* Texture filtering at camera-ray hits: `samples` rows are {raster x, raster y, lens x, lens y}. For each one the ray
 * differential of Camera::generateRayDifferential (src/librender/camera.cpp:79-90), narrowed by 1 / sqrt(sampleCount) as
 * SampleIntegrator::renderBlock does (src/librender/integrator.cpp:145,165), is intersected, the hit's BSDF is
 * fetched as Intersection::getBSDF(ray) does (include/mitsuba/render/records.inl:55-61, which runs
 * Intersection::computePartials, src/librender/intersection.cpp:5-75, when a texture filters), and `out` receives rows
 * of 12 floats: its.uv, dudx, dudy, dvdx, dvdy, then the BSDF's two spectrum slots as the textures return them
 * (Texture2D::getValue(its), src/librender/texture.cpp:73-82 -> MIPMap::getValue, src/librender/mipmap.cpp:264-316).
 * Zeros on a miss. */
MTSGPU_API int mtsgpu_texture_lookup(mtsgpu_ctx *ctx, const float *samples, size_t n, int sampleCount, float *out);

/* Closest-hit query that also fills the Intersection record (fillIntersectionRecord<true>,
 * skdtree.h:352-432): out rows of 12 floats {p, geoFrame.n, shFrame.n, wi}; zeros on a miss. */
MTSGPU_API int mtsgpu_intersect_full(mtsgpu_ctx *ctx, const float *rays, size_t n, mtsgpu_hit *hits, float *out);

/* The precomputed TriAccel records (48 B each, triaccel.h:34-46) in kd-tree primitive order
 * (shape-major), preceded by nothing; aabb6 receives the enlarged scene box (gkdtree.h:1172-1176). */
MTSGPU_API int mtsgpu_get_triaccel(mtsgpu_ctx *ctx, void *outRecords, size_t maxRecords, float aabb6[6]);

/* Acceleration-structure figures for the roofline accounting; the counterpart of the
 * "Structural kd-tree statistics" the reference logs after its build (include/mitsuba/render/gkdtree.h:1178-1214). */
typedef struct mtsgpu_accel_info {
	uint64_t triangles, nodes, nodeBytes, triBytes;
	double buildMs;
	uint32_t depth;        /* levels of the 4-wide tree; a traversal stack holds at most 3 entries per level */
	uint32_t builder;      /* 0 Morton radix tree, 1 PLOC, 2 binned SAH (the default) */
	uint32_t buildLevels;  /* levels of the binary SAH build (or PLOC rounds) */
	uint32_t reserved;
} mtsgpu_accel_info;
/* reads them back: what the reference prints as kd-tree statistics (include/mitsuba/render/gkdtree.h:1178-1214) */
MTSGPU_API int mtsgpu_accel_info_get(mtsgpu_ctx *ctx, mtsgpu_accel_info *info);

/* Per-ray traversal work (node visits, triangle tests) of the last mtsgpu_trace* call made with counters
 * enabled (mtsgpu_set_counters(ctx, 1)) — what rayIntersectHavranCollectStatistics returns per ray in the
 * reference (include/mitsuba/render/sahkdtree3.h:310-410). Off by default; a separate kernel instantiation,
 * so the timed kernels carry no counter code. */
typedef struct mtsgpu_trace_counters {
	uint64_t rays, nodeVisits, triTests;
	uint64_t maxStackEntries;   /* deepest traversal stack any ray reached since counters were switched on (trace calls and
	                             * renders; the limit is 96 entries, past which calls fail with MTSGPU_ERR_LIMIT) */
} mtsgpu_trace_counters;
MTSGPU_API int mtsgpu_set_counters(mtsgpu_ctx *ctx, int enabled);
MTSGPU_API int mtsgpu_get_counters(mtsgpu_ctx *ctx, mtsgpu_trace_counters *out);

#ifdef __cplusplus
}
#endif
#endif /* MTSGPU_H */
