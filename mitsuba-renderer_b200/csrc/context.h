// Host-side context of libmtscuda: owns the uploaded scene, the acceleration structure and
// the wavefront queues of one GPU. Internal; the C ABI in include/mtsgpu.h is the only surface.
#pragma once
#include <string>
#include <vector>
#include <cstdio>
#include <cstdarg>
#include <cstring>
#include <memory>
#include "types.cuh"

#define MTSGPU_PIPE_CHUNKS 16

struct DevBuf {
	void *p = nullptr; size_t bytes = 0;
	DevBuf() {}
	DevBuf(const DevBuf &) = delete;
	DevBuf &operator=(const DevBuf &) = delete;
	~DevBuf() { release(); }
	void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
	cudaError_t alloc(size_t n) {           // grow-only
		if (n <= bytes && p) return cudaSuccess;
		release();
		cudaError_t e = cudaMalloc(&p, n ? n : 16);
		if (e == cudaSuccess) bytes = n ? n : 16;
		return e;
	}
	template <typename T> T *as() const { return static_cast<T *>(p); }
};

// std::vector whose resize() leaves trivially constructible elements uninitialised: the mesh copies of a large scene are
// filled by a parallel memcpy right after (value-initialising 200 MB first costs as much as the copy)
template <typename T> struct NoInitAlloc : std::allocator<T> {
	template <typename U> struct rebind { using other = NoInitAlloc<U>; };
	NoInitAlloc() = default;
	template <typename U> NoInitAlloc(const NoInitAlloc<U> &) { }
	template <typename U> void construct(U *p) noexcept { ::new (static_cast<void *>(p)) U; }
	template <typename U, typename... A> void construct(U *p, A &&... a) { ::new (static_cast<void *>(p)) U(std::forward<A>(a)...); }
};
typedef std::vector<float, NoInitAlloc<float> > HostFloats;
typedef std::vector<uint32_t, NoInitAlloc<uint32_t> > HostIndices;

struct HostMesh {
	uint32_t nVert, nTri; int materialId, emitterId; bool hasNormals;
	HostFloats pos, nrm, uv; HostIndices tri;   // uv empty: no texture coordinates
	// analytic sphere (mtsgpu_upload_sphere): one vertex (the centre) and one primitive {0, SPHERE_MARK, SPHERE_MARK}
	bool sphere = false, inverted = false; float radius = 0, o2w[16] = {}, w2o[16] = {};
	// instance of a shape group (mtsgpu_upload_instance): two vertices (the corners of its world box, filled at build
	// time) and one primitive {0, INSTANCE_MARK, 1}; o2w / w2o as above
	bool instance = false; int groupId = -1;
	// analytic cylinder (mtsgpu_upload_cylinder): like an instance (two box corners, one primitive, o2w / w2o), radius above, length here
	bool cylinder = false; float length = 0;
};
#define INSTANCE_MARK 0xFFFFFFFEu
#define SPHERE_MARK 0xFFFFFFFFu

// Linear RGB pixels of one BITMAP texture (what MIPMap::fromBitmap receives), top row first
struct HostBitmap { int width = 0, height = 0; std::vector<float> rgb; };
// mipmap.cpp: the pyramid of MIPMap::MIPMap (src/librender/mipmap.cpp:31-137) appended to `texels` (rgb + pad per
// texel); `levels` receives {offset into texels, width, height, 0} per level
void buildMipPyramid(const HostBitmap &bmp, int filterType, int wrapMode, std::vector<float> &texels, std::vector<uint32_t> &levels);
void mipWeightLut(float *lut128);            // MIPMap's EWA weight table (mipmap.cpp:128-134)

struct mtsgpu_ctx {
	int device = 0;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev[8] = {};
	// upload / download streams + ordering events of the pipelined host-buffer ray query (api.cu)
	cudaStream_t copyIn = nullptr, copyOut = nullptr;
	cudaEvent_t pipeEv[2 * 16] = {};
	std::string error;

	// host-side scene staging
	std::vector<mtsgpu_material> materials;
	std::vector<mtsgpu_emitter> emitters;
	std::vector<mtsgpu_texture> textures;
	std::vector<HostBitmap> bitmaps;          // parallel to textures; filled by mtsgpu_upload_bitmap
	std::vector<float> emitterXf;             // 32 floats per emitter (luminaireToWorld, worldToLuminaire), zeros unless uploaded
	std::vector<mtsgpu_ctx *> groups;         // shape groups: child contexts with their own acceleration structure
	HostBitmap envBitmap; int envEmitter = -1; float envL2W[9] = {}, envW2L[9] = {};   // mtsgpu_upload_envmap
	uint32_t envTexture = 0;                  // index of the map's entry in the device texture table
	std::vector<HostMesh> meshes;
	std::vector<int> emitterShape;   // emitter id -> shape index
	int lumImportance = -1;          // which luminaire CDF is on the device (-1: none)
	CameraDev camera; bool haveCamera = false;
	FilmDev film; bool haveFilm = false;
	float filterTable[256];

	// device scene
	bool built = false;
	DeviceScene scene;
	DevBuf dTriAccel, dTriAccelOrig, dNodes, dTriIdx, dPositions, dNormals, dShapes, dMaterials, dEmitters,
	       dLumCdf, dAreaCdf, dTexcoords, dTextures, dTangents, dShadeTris, dSortedOf, dTexels, dMipLevels, dMipLut, dCamera, dEnvData, dInstances, dGroupScenes, dEmitterXf;
	mtsgpu_accel_info accel = {};

	// trace staging + counters
	DevBuf dRays, dHits, dCounters, dRayCounters, dTraceFlags;
	uint32_t rayCounterNext = 0;
	int smCount = 0;
	bool countersEnabled = false;
	mtsgpu_trace_counters lastCounters = {};

	// wavefront state (integrate.cu)
	DevBuf dPathState, dRayA, dRayB, dHit, dShadowRay, dShadowPayload, dQueueA, dQueueB, dQueueCounts,
	       dResult, dFilm, dStage, dStageOut;
	DevBuf dPrimSlot, dPrimRay, dPrimHit, dPrimState, dStateA, dStateB, dFilterTable;
	uint32_t *hCounts = nullptr;     // pinned host ring of polled queue lengths
	cudaEvent_t pollEv[16] = {};     // one per ring slot
	std::vector<cudaEvent_t> evPool[2];      // timing events of the traversal launches, two alternating sets
	DevBuf dWfStats;
	// progress / preview callback (mtsgpu_set_progress)
	mtsgpu_progress_fn progressFn = nullptr; void *progressUser = nullptr; double progressInterval = 0; bool progressWantFilm = false;
	cudaStream_t progressStream = nullptr; cudaEvent_t progressSnap = nullptr, progressCopied = nullptr;
	DevBuf dProgressSnap; float *hProgressFilm = nullptr; size_t hProgressBytes = 0;
	uint32_t pathCapacity = 0;
};

int setError(mtsgpu_ctx *ctx, int code, const char *fmt, ...);

#define CUDA_TRY(ctx, expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) \
	return setError((ctx), MTSGPU_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); } while (0)

#define REQUIRE(ctx, cond, code, ...) do { if (!(cond)) return setError((ctx), (code), __VA_ARGS__); } while (0)

// bvh_build.cu
int buildAccel(mtsgpu_ctx *ctx);
int rebuildLuminaireCdf(mtsgpu_ctx *ctx, bool importance);
// trace.cu
unsigned tracePersistentBlocks(mtsgpu_ctx *ctx, size_t n);
int traceRayCounter(mtsgpu_ctx *ctx, uint32_t **out);
int launchTrace(mtsgpu_ctx *ctx, const float4 *dRays, size_t n, int mode, void *dOut, uint64_t *dCounters);
int hostThreads();      // api.cu: threads of the host-side copy / flatten loops (MTSGPU_HOST_THREADS)
int checkTraceFlags(mtsgpu_ctx *ctx);
int traceRefetchBelow(const mtsgpu_ctx *ctx);
int traceLeafThreshold(const mtsgpu_ctx *ctx);
// integrate.cu
int renderDevice(mtsgpu_ctx *ctx, const mtsgpu_render_params *params, volatile const int *cancel,
	float *dOutFilm, mtsgpu_stats *stats);
int hookSamplerValues(mtsgpu_ctx *ctx, const mtsgpu_render_params *params, uint32_t pixel, uint32_t draw, float *out);
int hookBsdfEval(mtsgpu_ctx *ctx, const float *queries, size_t n, float *out);
int hookEmitterSample(mtsgpu_ctx *ctx, const float *queries, size_t n, float *out);
int hookCameraRays(mtsgpu_ctx *ctx, const float *samples, size_t n, float *out);
int hookFilmSplat(mtsgpu_ctx *ctx, const float *samples, size_t n, float *outFilm);
int hookIntersectUv(mtsgpu_ctx *ctx, const float *rays, size_t n, float *out);
int hookTextureLookup(mtsgpu_ctx *ctx, const float *samples, size_t n, int sampleCount, float *out);
int hookIntersectFull(mtsgpu_ctx *ctx, const float *rays, size_t n, mtsgpu_hit *hits, float *out);
