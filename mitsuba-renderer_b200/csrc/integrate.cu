// Wavefront path / direct-lighting integrator. Replaces, for one GPU and one sample range,
//   SampleIntegrator::render + BlockedRenderProcess + renderBlock   (src/librender/integrator.cpp:87-268,
//                                                                    src/librender/renderproc.cpp:106-160)
//   MIPathTracer::Li                                               (src/integrators/path/path.cpp:47-216)
//   MIDirectIntegrator::Li                                         (src/integrators/direct/direct.cpp:63-203)
//   ImageBlock::putSample + Film::putImageBlock                    (include/mitsuba/render/imageblock.h:80-138)
//
// The recursion of Li() becomes a loop over SoA queues in HBM:
//   k_generate        one camera ray per (pixel, sample) slot of the batch
//   k_closest_wf      closest hit for the active ray queue            (trace.cuh)
//   k_shade<MODE>     finish the previous bounce (MIS-weighted emitter hit, roulette, throughput),
//                     then at the new hit: emission, next-event estimation -> shadow queue,
//                     BSDF sampling -> next ray queue (warp-aggregated compaction)
//   k_shadow_wf       any-hit for the shadow queue; unoccluded contributions are added to the path
//   k_splat           per 16x16 tile: reconstruction-filter splat into shared memory, then one
//                     coalesced flush to the film accumulators {sum spec, sum alpha, sum weight}
// Batches are whole tiles x a chunk of samples, slot = (tile, sample, pixel-in-tile): a warp holds 32
// neighbouring pixels of one sample index, so primary rays are coherent and a warp's splats never
// collide on one pixel under the box filter.
#include "context.h"
#include "trace.cuh"
#include "bsdf.cuh"
#include "sampling.cuh"
#include <algorithm>
#include <chrono>
#include <cstdlib>

#define TILE 16
#define TILE_PIX (TILE * TILE)
#define WF_BLOCK 256
#define WF_POLL_RING 16           // pinned slots / events of the asynchronous queue-length polls
#define WF_POLL_AHEAD 6           // bounces the host may enqueue beyond the last poll it has seen
#define SHADE_CHUNK 1024          // queue entries a shading block lists before it shades the ones that need it

// path-state flag word
#define PF_DEPTH_MASK   0x0000FFFFu
#define PF_EMITTED      0x00010000u   // EEmittedRadiance still set (camera hit only, path.cpp:191)
#define PF_PREV_DELTA   0x00020000u   // previous BSDF sample was a delta lobe
#define PF_PREV_TRANS   0x00040000u   // previous BSDF sample was transmission
#define PF_PENDING      0x00080000u   // a BSDF-sampled ray is in flight (steps 7-8 of the loop are outstanding)
#define PF_RR_SHIFT     20            // bits 20-27: roulette values drawn so far (the halton sampler's value counter needs it)
#define PF_RR_MASK      0x0FF00000u

struct WfArrays {
	float4 *li;         // per slot: Li.rgb, alpha — touched only when a path adds radiance (emitter hit, background, unoccluded shadow ray)
	// per queue position, travelling with the ray (2 x float4: {throughput.rgb, previous bsdfPdf} {previous bsdfVal.rgb (f*cos/pdf),
	// flag word}): read and written in queue order, where the slot-indexed arrays of the first version cost six random
	// 32-byte sectors per hit from the second bounce on (k_shade was bound by DRAM transactions, ~50 G sectors/s)
	float4 *stateCur, *stateNext, *statePrim;
	float2 *samplePos;  // per slot: film sample position; x = NaN marks a slot outside the crop window
	uint32_t *slotCur, *slotNext;     // ray queues: slot ids by queue position
	float4 *rayCur, *rayNext;         // 2 x float4 per queue position
	uint4 *hit;                       // per queue position {t, u, v, sorted triangle}
	float4 *shadowRay;                // 2 x float4 per shadow-queue position
	float4 *shadowPayload;            // contribution.rgb, slot
	uint32_t *counters;               // [0] next-queue count, [1] shadow-queue count
	// direct mode keeps the primary hit while secondary rays are traced
	uint32_t *slotPrim; float4 *rayPrim; uint4 *hitPrim;
	uint32_t capacity;                // the hit arrays are followed by `capacity` nested-triangle indices (instance hits)
};

struct BatchDesc {
	uint32_t tileBegin, tileCount;       // tiles of this batch (row-major tile index over the crop window)
	uint32_t tilesX;
	uint32_t sampleBegin, sampleCount;   // global sample indices [sampleBegin, sampleBegin + sampleCount)
	uint32_t slots;                      // tileCount * sampleCount * TILE_PIX
	uint32_t seed;
	float diffScale;                     // 1 / sqrt(sample count): SampleIntegrator::renderBlock narrows the ray differentials by it
	SamplerConfig sampler;
};

struct IntegratorDev {
	int mode, maxDepth, rrDepth, strictNormals, nLum, nBsdf;
	int importance;
};

namespace {

__device__ __forceinline__ uint32_t warpAppend(uint32_t *counter, bool pred) {
	// warp-aggregated queue append: one atomic per warp, ballot + prefix popcount for the lane offset
	const uint32_t mask = __ballot_sync(0xFFFFFFFFu, pred);
	if (mask == 0) return 0;
	const int lane = threadIdx.x & 31;
	const int leader = __ffs(mask) - 1;
	uint32_t base = 0;
	if (lane == leader) base = atomicAdd(counter, (uint32_t) __popc(mask));
	base = __shfl_sync(0xFFFFFFFFu, base, leader);
	return base + __popc(mask & ((1u << lane) - 1u));
}

// Block-aggregated queue append: ONE atomic per block and call. Appends to one address serialise in L2 at about 145 M/s on
// B200 (k_generate: 4.2 M warp-level appends took 29 ms), and with one append per warp k_shade needed ~125 M/s on each of
// its two queues: it was bound by exactly that. `agg` / `base`: shared scratch, double-buffered by `parity` so that one
// barrier pair per call suffices. Must be called by every thread of the block.
__device__ __forceinline__ uint32_t blockAppend(uint32_t *counter, bool pred, uint32_t (*agg)[WF_BLOCK / 32], uint32_t *base, int parity) {
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint32_t mask = __ballot_sync(0xFFFFFFFFu, pred);
	if (lane == 0) agg[parity][warp] = __popc(mask);
	__syncthreads();
	if (threadIdx.x == 0) {
		uint32_t total = 0;
#pragma unroll
		for (int w = 0; w < WF_BLOCK / 32; ++w) total += agg[parity][w];
		base[parity] = total ? atomicAdd(counter, total) : 0u;
	}
	__syncthreads();
	uint32_t pos = base[parity];
	for (uint32_t w = 0; w < warp; ++w) pos += agg[parity][w];
	return pos + __popc(mask & ((1u << lane) - 1u));
}

// Two queues at once: the counters of the next-ray and the shadow queue are the halves of one aligned 64-bit word, so a
// block reserves room in both with ONE atomic behind ONE barrier pair. With an append per queue k_shade's blocks sat at
// barriers for two atomic round trips per 256 entries (ncu, profiles/r02_final_c4_bounce2: barrier = 2.8 of 13.7 stall
// cycles per issue). Per-warp counts travel packed (16 bits each: a block emits at most WF_BLOCK entries per queue and call).
__device__ __forceinline__ void blockAppend2(unsigned long long *counter2, bool predA, bool predB, uint32_t (*agg)[WF_BLOCK / 32],
		unsigned long long *base, int parity, uint32_t &posA, uint32_t &posB) {
	static_assert(WF_BLOCK < 65536, "packed per-warp counts");
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint32_t maskA = __ballot_sync(0xFFFFFFFFu, predA), maskB = __ballot_sync(0xFFFFFFFFu, predB);
	if (lane == 0) agg[parity][warp] = __popc(maskA) | (__popc(maskB) << 16);
	__syncthreads();
	if (threadIdx.x == 0) {
		uint32_t total = 0;
#pragma unroll
		for (int w = 0; w < WF_BLOCK / 32; ++w) total += agg[parity][w];
		base[parity] = total ? atomicAdd(counter2, (unsigned long long) (total & 0xFFFFu) | ((unsigned long long) (total >> 16) << 32)) : 0ull;
	}
	__syncthreads();
	const unsigned long long b = base[parity];
	uint32_t before = 0;
	for (uint32_t w = 0; w < warp; ++w) before += agg[parity][w];
	const uint32_t lt = (1u << lane) - 1u;
	posA = (uint32_t) b + (before & 0xFFFFu) + __popc(maskA & lt);
	posB = (uint32_t) (b >> 32) + (before >> 16) + __popc(maskB & lt);
}

__device__ __forceinline__ void slotToPixel(const BatchDesc &b, const FilmDev &film, uint32_t slot,
		int &px, int &py, uint32_t &sample) {
	const uint32_t pix = slot % TILE_PIX;
	const uint32_t rest = slot / TILE_PIX;
	const uint32_t s = rest % b.sampleCount;
	const uint32_t tile = b.tileBegin + rest / b.sampleCount;
	px = (int) ((tile % b.tilesX) * TILE + pix % TILE);
	py = (int) ((tile / b.tilesX) * TILE + pix / TILE);
	sample = b.sampleBegin + s;
}

__global__ void __launch_bounds__(WF_BLOCK)
k_generate(const CameraDev cam, const FilmDev film, const BatchDesc b, WfArrays wf) {
	const uint32_t slot = blockIdx.x * WF_BLOCK + threadIdx.x;
	bool alive = false;
	float4 r0, r1;
	if (slot < b.slots) {
		int px, py; uint32_t sample;
		slotToPixel(b, film, slot, px, py, sample);
		if (px < film.cropW && py < film.cropH) {
			alive = true;
			const int fx = px + film.cropX, fy = py + film.cropY;     // sensor pixel
			Sampler smp; smp.init((uint32_t) (fy * film.width + fx), sample, b.seed, b.sampler);
			float lx = 0, ly = 0, sx, sy;
			if (cam.apertureRadius > 0.0f) smp.get2D(DIM_LENS, lx, ly);      // renderBlock: lens, time, then pixel sample
			smp.get2D(DIM_PIXEL, sx, sy);
			sx += (float) fx; sy += (float) fy;
			V3 o, d; float mint, maxt;
			cameraRay(cam, sx, sy, lx, ly, o, d, mint, maxt);
			r0 = make_float4(o.x, o.y, o.z, mint);
			r1 = make_float4(d.x, d.y, d.z, maxt);
			wf.samplePos[slot] = make_float2(sx, sy);
			wf.li[slot] = make_float4(0, 0, 0, 0);
		} else {
			wf.samplePos[slot] = make_float2(__int_as_float(0x7FC00000), 0);
		}
	}
	__shared__ uint32_t sAgg[2][WF_BLOCK / 32], sBase[2];
	const uint32_t pos = blockAppend(&wf.counters[0], alive, sAgg, sBase, 0);
	if (alive) {
		wf.slotNext[pos] = slot;
		wf.rayNext[2 * (size_t) pos] = r0;
		wf.rayNext[2 * (size_t) pos + 1] = r1;
		wf.stateNext[2 * (size_t) pos] = make_float4(1, 1, 1, 0);
		wf.stateNext[2 * (size_t) pos + 1] = make_float4(0, 0, 0, __uint_as_float(1u | PF_EMITTED));   // depth = 1 (integrator.h:186-191)
	}
}

// Wavefront traversal kernels, one thread per queue entry (lock-step). Measured inside renders at 64 spp
// (B200): closest-hit queues run faster through the persistent-warp variant below (C3 33.2 -> 25.6 ms,
// C4 177.8 -> 167.1 ms), shadow queues — coherent rays towards a handful of emitters — do not on the
// 10M-triangle scene (61.0 -> 71.3 ms), so shadow rays keep this kernel. MTSGPU_WF_PERSIST_{CLOSEST,SHADOW}
// override the choice for experiments.
template <bool SPH>
__global__ void __launch_bounds__(TRACE_BLOCK)
k_closest_wf(const DeviceScene sc, const float4 *__restrict__ rays, const uint32_t *__restrict__ nPtr, uint4 *__restrict__ hits, uint32_t *__restrict__ nestedOut) {
	__shared__ uint2 sstack[TRACE_SMEM_STACK * TRACE_BLOCK];
	const uint32_t n = *nPtr;                  // the queue length lives on the device: no host round trip between bounces
	for (uint32_t i = blockIdx.x * TRACE_BLOCK + threadIdx.x; i < n; i += gridDim.x * TRACE_BLOCK) {
		const float4 a = __ldg(rays + 2 * (size_t) i), b = __ldg(rays + 2 * (size_t) i + 1);
		RayIn r = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
		HitOut ho; ho.t = INFINITY; ho.u = ho.v = 0; ho.tri = 0xFFFFFFFFu; ho.nested = 0xFFFFFFFFu;
		float mint, maxt; uint32_t c0 = 0, c1 = 0;
		if (clipToScene<false>(sc, r, mint, maxt))
			traverse<false, false, SPH>(sc, r, mint, maxt, sstack, ho, c0, c1);
		hits[i] = make_uint4(__float_as_uint(ho.t), __float_as_uint(ho.u), __float_as_uint(ho.v), ho.tri);
		if (SPH && nestedOut) nestedOut[i] = ho.nested;
	}
}

template <bool SPH>
__global__ void __launch_bounds__(TRACE_BLOCK)
k_shadow_wf(const DeviceScene sc, const float4 *__restrict__ rays, const float4 *__restrict__ payload, const uint32_t *__restrict__ nPtr, float4 *li) {
	__shared__ uint2 sstack[TRACE_SMEM_STACK * TRACE_BLOCK];
	const uint32_t n = *nPtr;
	for (uint32_t i = blockIdx.x * TRACE_BLOCK + threadIdx.x; i < n; i += gridDim.x * TRACE_BLOCK) {
		const float4 a = __ldg(rays + 2 * (size_t) i), b = __ldg(rays + 2 * (size_t) i + 1);
		RayIn r = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
		float mint, maxt; uint32_t c0 = 0, c1 = 0;
		bool occluded = false;
		if (clipToScene<true>(sc, r, mint, maxt)) {
			HitOut ho;
			occluded = traverse<true, false, SPH>(sc, r, mint, maxt, sstack, ho, c0, c1);
		}
		if (!occluded) {
			const float4 p = __ldg(payload + i);
			const uint32_t slot = __float_as_uint(p.w);
			float4 l = li[slot];                 // one shadow ray per slot and pass: plain read-modify-write
			l.x += p.x; l.y += p.y; l.z += p.z;
			li[slot] = l;
		}
	}
}

// Persistent-warp variants (trace.cuh) of the two wavefront traversal kernels.
struct WfHitSink {
	uint4 *hits;
	__device__ __forceinline__ void closest(uint32_t ray, bool found, float t, float u, float v, uint32_t tri) {
		hits[ray] = make_uint4(__float_as_uint(found ? t : INFINITY), __float_as_uint(u), __float_as_uint(v), found ? tri : 0xFFFFFFFFu);
	}
	__device__ __forceinline__ void any(uint32_t, bool) { }
};
struct WfShadowSink {
	const float4 *payload; float4 *li;
	__device__ __forceinline__ void closest(uint32_t, bool, float, float, float, uint32_t) { }
	__device__ __forceinline__ void any(uint32_t ray, bool occluded) {
		if (occluded) return;
		const float4 p = __ldg(payload + ray);
		const uint32_t slot = __float_as_uint(p.w);
		float4 l = li[slot];
		l.x += p.x; l.y += p.y; l.z += p.z;
		li[slot] = l;
	}
};
// COUNT: the counter build (mtsgpu_set_counters): node visits and primitive tests of the render's own ray queues, summed into
// `counters` — a separate instantiation, so the timed kernels carry no counter code.
template <bool SPH, bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK, TRACE_PERSISTENT_BLOCKS_PER_SM)
k_closest_wfp(const DeviceScene sc, const float4 *__restrict__ rays, const uint32_t *__restrict__ nPtr, uint4 *__restrict__ hits, uint32_t *rayCounter, int refetchBelow, int leafThreshold,
		unsigned long long *counters) {
	__shared__ uint2 sstack[TRACE_SMEM_STACK * TRACE_BLOCK];
	uint32_t c0 = 0, c1 = 0;
	const uint32_t n = *nPtr;
	WfHitSink sink = { hits };
	tracePersistent<false, COUNT, WfHitSink, SPH>(sc, rays, n, rayCounter, sink, sstack, c0, c1, refetchBelow, true, leafThreshold);
	if (COUNT) flushCounters(c0, c1, counters);
}
template <bool SPH, bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK, TRACE_PERSISTENT_BLOCKS_PER_SM)
k_shadow_wfp(const DeviceScene sc, const float4 *__restrict__ rays, const float4 *__restrict__ payload, const uint32_t *__restrict__ nPtr, float4 *li, uint32_t *rayCounter, int refetchBelow, int leafThreshold,
		unsigned long long *counters) {
	__shared__ uint2 sstack[TRACE_SMEM_STACK * TRACE_BLOCK];
	uint32_t c0 = 0, c1 = 0;
	const uint32_t n = *nPtr;
	WfShadowSink sink = { payload, li };
	tracePersistent<true, COUNT, WfShadowSink, SPH>(sc, rays, n, rayCounter, sink, sstack, c0, c1, refetchBelow, true, leafThreshold);
	if (COUNT) flushCounters(c0, c1, counters);
}

__device__ __forceinline__ float miWeight(float pdfA, float pdfB) {     // path.cpp:218-222 (power heuristic)
	pdfA *= pdfA; pdfB *= pdfB;
	return pdfA / (pdfA + pdfB);
}

// Next-event estimation at `its` (path.cpp:97-125 / direct.cpp:128-150). Returns true and fills the
// shadow ray + contribution when a visibility test is needed.
__device__ __forceinline__ bool nextEvent(const DeviceScene &sc, const IntegratorDev &ig, const Isect &its,
		const MaterialDev &mat, V3 rayD, V3 scale, float sx, float sy, float fracLum, float fracBsdf, float weightLum,
		float4 &s0, float4 &s1, V3 &contrib) {
	LumSample ls;
	sampleLuminaire(sc, its.p, sx, sy, ig.importance != 0, ls);
	if (ls.pdf == 0) return false;
	const V3 wo = -ls.d;
	const V3 woLocal = toLocal(its.sh, wo);
	V3 f; float bPdf;
	bsdfFPdf(mat, its.wi, woLocal, f, bPdf);
	const V3 bsdfVal = f * fabsf(woLocal.z);                               // fCos
	if (isZero(bsdfVal)) return false;
	if (!ls.mis) bPdf = 0.0f;                                              // delta luminaires: BSDF sampling cannot find them
	if (isZero(ls.value)) return false;                                    // outside a spot's cone: nothing to add, no shadow ray needed
	if (ig.strictNormals && !(dot(its.geoN, wo) * woLocal.z > 0)) return false;
	const float weight = miWeight(ls.pdf * fracLum, bPdf * fracBsdf) * weightLum;
	contrib = ((scale * ls.value) * bsdfVal) * weight;
	// Scene::isOccluded(p1, p2): Ray(p1, p2 - p1, ShadowEpsilon, 1 - ShadowEpsilon) (scene.h:241-246)
	const V3 dd = ls.p - its.p;
	s0 = make_float4(its.p.x, its.p.y, its.p.z, MTS_SHADOW_EPSILON);
	s1 = make_float4(dd.x, dd.y, dd.z, 1.0f - MTS_SHADOW_EPSILON);
	return true;
}

// One shading step for the queue entry i. MODE 0 = path, 1 = direct (primary pass), 2 = direct (secondary pass).
// BMP: the scene has bitmap textures or instances (the variant without them carries none of that code).
#ifndef SHADE_BINS
#define SHADE_BINS 4      // BSDF classes k_shade orders its list by (1: queue order only)
#endif
// Class of the BSDF a hit will run: 0 ideally diffuse or none, 1 delta (mirror, dielectric), 2 Beckmann microfacet (roughmetal,
// microfacet), 3 the rest (phong, ward, roughglass). Only the order of the work depends on it.
__device__ __forceinline__ uint32_t shadeClass(const DeviceScene &sc, uint32_t tri) {
	const uint32_t shape = __float_as_uint(__ldg(&sc.shadeTris[6 * (size_t) tri].w));
	const int mat = sc.shapes[shape].materialId;
	if (mat < 0) return 0u;
	const uint32_t t = sc.materials[mat].type & MTSGPU_BSDF_TYPE_MASK;
	if (t == MTSGPU_BSDF_ROUGHMETAL || t == MTSGPU_BSDF_MICROFACET) return 2u % SHADE_BINS;
	if (t == MTSGPU_BSDF_MIRROR || t == MTSGPU_BSDF_DIELECTRIC) return 1u % SHADE_BINS;
	if (t == MTSGPU_BSDF_LAMBERTIAN || t == MTSGPU_BSDF_DIFFTRANS) return 0u;
	return 3u % SHADE_BINS;
}
template <int MODE, bool BMP>
#ifndef SHADE_MIN_BLOCKS
#define SHADE_MIN_BLOCKS 3
#endif
__global__ void __launch_bounds__(WF_BLOCK, SHADE_MIN_BLOCKS)      // 3 x 256 threads: <= 85 registers
k_shade(const DeviceScene sc, const IntegratorDev ig, const BatchDesc b, const FilmDev film, WfArrays wf, const uint32_t *__restrict__ nIn, uint32_t round) {
	// The queue length is read on the device (no host round trip per bounce), and a block works through chunks of
	// SHADE_CHUNK queue entries: it first lists the entries that need shading at all — from the second bounce on more than
	// half of the rays have left the scene, and a miss without a background luminaire changes nothing (ncu, r02_shade_base:
	// 9.9 of 32 lanes active on the bounce-2 queue) — then shades the listed ones with full warps. The list keeps the queue order.
	__shared__ uint32_t sList[SHADE_CHUNK];
	__shared__ uint32_t sCount[SHADE_BINS * (SHADE_CHUNK / 32)];      // per (class, warp-sized piece of the chunk): entries, then their offset in the list
	__shared__ uint32_t sListed;
	__shared__ uint32_t sAgg[2][WF_BLOCK / 32];
	__shared__ unsigned long long sBase[2];
	int parity = 0;
	const uint32_t n = *nIn;
	const uint32_t *slots = (MODE == 1) ? wf.slotPrim : wf.slotCur;
	const float4 *rays = (MODE == 1) ? wf.rayPrim : wf.rayCur;
	const uint4 *hits = (MODE == 1) ? wf.hitPrim : wf.hit;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	for (uint32_t chunk0 = blockIdx.x * SHADE_CHUNK; chunk0 < n; chunk0 += gridDim.x * SHADE_CHUNK) {
	// The list is ordered by BSDF class, queue order within a class (a counting sort over the chunk): the classes differ in the
	// code they run — ncu on the C4 queues: the rough-metal sampling code ran with 13.6 of 31 lanes, the other lanes held hits on
	// the glass spheres — so warps that see one class run it with full lanes. The class of a hit costs one sector of its shading
	// record (fetched again, from cache, by the shading pass) and two table lookups.
	uint32_t ownMask[SHADE_CHUNK / WF_BLOCK], ownBin[SHADE_CHUNK / WF_BLOCK];
#pragma unroll
	for (int k = 0; k < SHADE_CHUNK / WF_BLOCK; ++k) {
		const uint32_t idx = chunk0 + k * WF_BLOCK + threadIdx.x;
		const uint32_t tri = idx < n ? __ldg(&hits[idx].w) : 0xFFFFFFFFu;
		const bool need = idx < n && (sc.background >= 0 || tri != 0xFFFFFFFFu);
		uint32_t bin = 0;
		if (SHADE_BINS > 1 && tri != 0xFFFFFFFFu) bin = shadeClass(sc, tri);
		uint32_t mine = 0;
#pragma unroll
		for (uint32_t bb = 0; bb < SHADE_BINS; ++bb) {
			const uint32_t m = __ballot_sync(0xFFFFFFFFu, need && bin == bb);
			if (lane == 0) sCount[bb * (SHADE_CHUNK / 32) + k * (WF_BLOCK / 32) + warp] = __popc(m);
			if (bin == bb) mine = m;
		}
		ownMask[k] = need ? mine : 0u; ownBin[k] = bin;
	}
	__syncthreads();
	if (warp == 0) {
		// exclusive scan of the SHADE_BINS * 32 counts: every lane sums its SHADE_BINS consecutive counts, the lanes scan by shuffles
		uint32_t c[SHADE_BINS], sum = 0;
#pragma unroll
		for (int q = 0; q < SHADE_BINS; ++q) { c[q] = sCount[lane * SHADE_BINS + q]; sum += c[q]; }
		uint32_t incl = sum;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o); if ((int) lane >= o) incl += v; }
		uint32_t run = incl - sum;
#pragma unroll
		for (int q = 0; q < SHADE_BINS; ++q) { sCount[lane * SHADE_BINS + q] = run; run += c[q]; }
		if (lane == 31) sListed = incl;
	}
	__syncthreads();
#pragma unroll
	for (int k = 0; k < SHADE_CHUNK / WF_BLOCK; ++k)
		if (ownMask[k])
			sList[sCount[ownBin[k] * (SHADE_CHUNK / 32) + k * (WF_BLOCK / 32) + warp] + __popc(ownMask[k] & ((1u << lane) - 1u))] = chunk0 + k * WF_BLOCK + threadIdx.x;
	__syncthreads();
	const uint32_t listed = sListed;
	for (uint32_t j0 = 0; j0 < listed; j0 += WF_BLOCK) {
	const bool entry = j0 + threadIdx.x < listed;
	const uint32_t i = entry ? sList[j0 + threadIdx.x] : 0u;
	bool emitRay = false, emitShadow = false;
	float4 nr0, nr1, sr0, sr1, st0, st1; V3 contrib = v3(0, 0, 0);
	uint32_t slot = 0;

	if (entry) {
		slot = slots[i];
		const float4 ra = rays[2 * (size_t) i], rb = rays[2 * (size_t) i + 1];
		const uint4 h = hits[i];
		const V3 rayO = v3(ra), rayD = v3(rb);
		const bool valid = h.w != 0xFFFFFFFFu;
		const float4 *state = (MODE == 1) ? wf.statePrim : wf.stateCur;
		float4 thr4 = state[2 * (size_t) i], bv4 = state[2 * (size_t) i + 1];
		uint32_t flags = __float_as_uint(bv4.w);
		uint32_t depth = flags & PF_DEPTH_MASK;
		V3 thr = v3(thr4), Li = v3(0, 0, 0);            // Li: the radiance this step adds (at most one term), folded into li[slot] at the end
		bool liDirty = false, alphaDirty = false;
		int px, py; uint32_t sample;
		slotToPixel(b, film, slot, px, py, sample);
		Sampler smp; smp.init((uint32_t) ((py + film.cropY) * film.width + px + film.cropX), sample, b.seed, b.sampler);
		smp.rouletteDraws = (flags & PF_RR_MASK) >> PF_RR_SHIFT;

		Isect its;
		bool alive = valid;
		if (valid) {
			uint32_t nested = 0xFFFFFFFFu;
			if (BMP && sc.hasInstances) nested = reinterpret_cast<const uint32_t *>(hits + wf.capacity)[i];
			fillIntersectionSorted<BMP>(sc, h.w, __uint_as_float(h.y), __uint_as_float(h.z), rayO, rayD, __uint_as_float(h.x), its, nested);
		}

		if (MODE == 2) {
			// direct.cpp:170-199: BSDF-sampled ray — only emitters (or the background on a miss) contribute
			if ((valid && its.emitterId >= 0) || (!valid && sc.background >= 0)) {
				const V3 Le = valid ? emitterLe(sc, its.emitterId, its.geoN, -rayD) : backgroundLe(sc, rayD);
				const float lumPdf = (flags & PF_PREV_DELTA) ? 0.0f :
					(valid ? pdfLuminaire(sc, rayO, its.p, its.geoN, its.emitterId, ig.importance != 0) : pdfBackground(sc, ig.importance != 0, rayD));
				const float total = (float) (ig.nLum + ig.nBsdf);          // direct.cpp:49-54
				const float fracBsdf = ig.nBsdf / total, fracLum = ig.nLum / total;
				const float weightBsdf = 1 / (float) ig.nBsdf;
				const float w = miWeight(thr4.w * fracBsdf, lumPdf * fracLum) * weightBsdf;
				Li = Li + (Le * v3(bv4)) * w; liDirty = true;
			}
			alive = false;
		} else {
			if (MODE == 0 && (flags & PF_PENDING)) {
				// ---- steps after "trace a ray in this direction" of the previous iteration (path.cpp:152-208)
				const V3 bsdfVal = v3(bv4);
				if (!valid) {
					if (sc.background >= 0) {                  // path.cpp:158-163, 172-179, then 187-188
						const float lumPdf = (flags & PF_PREV_DELTA) ? 0.0f : pdfBackground(sc, ig.importance != 0, rayD);
						const float w = miWeight(thr4.w, lumPdf);
						Li = Li + ((thr * backgroundLe(sc, rayD)) * bsdfVal) * w; liDirty = true;
					} else {
						depth++;                               // path.cpp:166
					}
					alive = false;
				} else {
					if (its.emitterId >= 0) {
						const V3 Le = emitterLe(sc, its.emitterId, its.geoN, -rayD);
						const float lumPdf = (flags & PF_PREV_DELTA) ? 0.0f : pdfLuminaire(sc, rayO, its.p, its.geoN, its.emitterId, ig.importance != 0);
						const float w = miWeight(thr4.w, lumPdf);
						Li = Li + ((thr * Le) * bsdfVal) * w; liDirty = true;
					}
					flags &= ~PF_EMITTED;                      // rRec.type = ERadianceNoEmission
					bool survive = true;
					if ((int) depth >= ig.rrDepth && !(flags & PF_PREV_TRANS)) {
						const float approxAlbedo = fminf(0.9f, maxc(bsdfVal));
						const float rr = smp.get1D(DIM_BOUNCE(depth - 1, 2), smp.rouletteDraws);      // the path's k-th roulette draw is its k-th next1D() (after the time sample)
						smp.rouletteDraws = min(smp.rouletteDraws + 1u, 255u);
						flags = (flags & ~PF_RR_MASK) | (smp.rouletteDraws << PF_RR_SHIFT);
						if (rr > approxAlbedo) survive = false;
						else thr = thr * (1.0f / approxAlbedo);
					}
					if (survive) { thr = thr * bsdfVal; depth++; }
					else alive = false;
				}
			} else if (MODE == 0 || round == 0) {
				// camera hit: rRec.alpha (records.inl:93-99)
				alphaDirty = valid;               // k_generate left alpha = 0
				if (MODE == 0 && ig.maxDepth == 0) alive = false;      // path.cpp:60: `while (depth <= maxDepth || maxDepth < 0)` never runs: black
				else if (!valid && (flags & PF_EMITTED)) { Li = Li + backgroundLe(sc, rayD); liDirty = true; }      // path.cpp:62-68 / direct.cpp:74-81
			}

			// ---- top of the loop body at the new hit (path.cpp:58-95 / direct.cpp:74-100)
			if (alive && its.materialId < 0 && its.emitterId >= 0 && (flags & PF_EMITTED) && (MODE == 0 || round == 0)) {
				// direct.cpp adds emission before the BSDF check; path.cpp checks the BSDF first
				if (MODE == 1) { Li = Li + emitterLe(sc, its.emitterId, its.geoN, -rayD); liDirty = true; }
			}
			if (alive && its.materialId < 0) alive = false;             // no BSDF: give up (path.cpp:70-77)
			MaterialDev mat;
			if (alive) {
				mat = sc.materials[its.materialId];
				float4 duv = make_float4(0, 0, 0, 0);
				if (BMP && MODE != 2 && depth == 1 && (mat.tex[0] | mat.tex[1] | mat.tex[2]) && usesRayDifferentials(sc, mat)) {
					// Intersection::getBSDF(ray) (records.inl:55-61): the camera ray is the only one with differentials
					float lx = 0, ly = 0;
					if (b.sampler.lensDraw) smp.get2D(DIM_LENS, lx, ly);
					const float2 sp = wf.samplePos[slot];
					duv = cameraHitPartials(sc, sp.x, sp.y, lx, ly, rayO, rayD, b.diffScale, h.w, __uint_as_float(h.y), __uint_as_float(h.z), its.p, its.geoN,
						sc.hasInstances ? reinterpret_cast<const uint32_t *>(hits + wf.capacity)[i] : 0xFFFFFFFFu);
				}
				resolveTextures<BMP>(sc, mat, its.uvx, its.uvy, duv);
				if (its.emitterId >= 0 && (flags & PF_EMITTED) && (MODE == 0 || round == 0))
					{ Li = Li + (MODE == 0 ? thr : v3(1, 1, 1)) * emitterLe(sc, its.emitterId, its.geoN, -rayD); liDirty = true; }
				if (MODE == 0 && ig.maxDepth > 0 && (int) depth >= ig.maxDepth) alive = false;
				if (MODE == 0 && depth >= PF_DEPTH_MASK - 1u) alive = false;      // safety net: the depth field is 16 bits wide
			}
			if (alive && MODE == 0 && ig.strictNormals) {
				const float wiDotGeoN = -dot(its.geoN, rayD);
				if (wiDotGeoN * its.wi.z < 0) alive = false;
			}
			if (alive) {
				float fracLum = 1.0f, fracBsdf = 1.0f, weightLum = 1.0f, weightBsdf = 1.0f;
				bool doLum = true, doBsdf = true;
				uint32_t dimL, dimB;
				if (MODE == 1) {
					const float total = (float) (ig.nLum + ig.nBsdf);
					fracLum = ig.nLum / total; fracBsdf = ig.nBsdf / total;
					weightLum = 1 / (float) ig.nLum; weightBsdf = 1 / (float) ig.nBsdf;
					doLum = (int) round < ig.nLum; doBsdf = (int) round < ig.nBsdf;
					dimL = DIM_DIRECT(round, 0); dimB = DIM_DIRECT(round, 1);
				} else {
					dimL = DIM_BOUNCE(depth - 1, 0); dimB = DIM_BOUNCE(depth - 1, 1);
				}
				(void) weightBsdf;
				// ---- luminaire sampling
				if (doLum) {
					float sx, sy; smp.get2D(dimL, sx, sy);
					// delta-only BSDFs have f = 0: the reference still traces the shadow ray, the result is unused
					if ((mat.type & MTSGPU_BSDF_TYPE_MASK) != MTSGPU_BSDF_MIRROR && (mat.type & MTSGPU_BSDF_TYPE_MASK) != MTSGPU_BSDF_DIELECTRIC)
						emitShadow = nextEvent(sc, ig, its, mat, rayD, MODE == 0 ? thr : v3(1, 1, 1), sx, sy,
							fracLum, fracBsdf, weightLum, sr0, sr1, contrib);
				}
				// ---- BSDF sampling
				bool cont = false;
				if (doBsdf) {
					float sx, sy; smp.get2D(dimB, sx, sy);
					V3 woLocal; float pdf = 0; uint32_t type = 0;
					V3 val = bsdfSample3(mat, its.wi, sx, sy, woLocal, pdf, type);
					if (!isZero(val)) {
						val = val * fabsf(woLocal.z);                         // sampleCos
						if (!isZero(val)) {
							val = val * (1.0f / pdf);                         // bsdfVal /= bsdfPdf
							const V3 wo = toWorld(its.sh, woLocal);
							bool ok = true;
							if (MODE == 0 && ig.strictNormals && dot(its.geoN, wo) * woLocal.z <= 0) ok = false;
							if (ok) {
								cont = true;
								nr0 = make_float4(its.p.x, its.p.y, its.p.z, MTS_EPSILON);
								nr1 = make_float4(wo.x, wo.y, wo.z, INFINITY);
								flags &= ~(PF_PREV_DELTA | PF_PREV_TRANS);
								if (type & BSDF_DELTA) flags |= PF_PREV_DELTA;
								if (type & BSDF_TRANSMISSION) flags |= PF_PREV_TRANS;
								flags |= PF_PENDING;
								bv4 = make_float4(val.x, val.y, val.z, 0);
								thr4.w = pdf;
							}
						}
					}
				}
				emitRay = cont;
				if (MODE == 0 && !cont) alive = false;
			}
			flags = (flags & ~PF_DEPTH_MASK) | (depth & PF_DEPTH_MASK);
			if (smp.overflow) atomicOr(sc.traceFlags, 2u);      // halton.cpp:74-84: the reference stops with an error here
			if (emitRay) { thr4 = make_float4(thr.x, thr.y, thr.z, thr4.w); bv4.w = __uint_as_float(flags); }
		}
		if (liDirty || alphaDirty) {
			float4 l = wf.li[slot];
			if (liDirty) { l.x += Li.x; l.y += Li.y; l.z += Li.z; }
			if (alphaDirty) l.w = 1.0f;
			wf.li[slot] = l;
		}
		if (emitRay) { st0 = thr4; st1 = bv4; }
	}

	uint32_t rp, sp;
	blockAppend2(reinterpret_cast<unsigned long long *>(&wf.counters[2]), emitRay, emitShadow, sAgg, sBase, parity, rp, sp);
	parity ^= 1;
	if (emitRay) {
		wf.slotNext[rp] = slot;
		wf.rayNext[2 * (size_t) rp] = nr0;
		wf.rayNext[2 * (size_t) rp + 1] = nr1;
		wf.stateNext[2 * (size_t) rp] = st0;
		wf.stateNext[2 * (size_t) rp + 1] = st1;
	}
	if (emitShadow) {
		wf.shadowRay[2 * (size_t) sp] = sr0;
		wf.shadowRay[2 * (size_t) sp + 1] = sr1;
		wf.shadowPayload[sp] = make_float4(contrib.x, contrib.y, contrib.z, __uint_as_float(slot));
	}
	}      // listed entries
	__syncthreads();      // sList is rewritten by the next chunk
	}      // chunks
}

// End of a bounce: add the queue lengths to the statistics, make the next ray queue the current one, empty the others.
// counters: [0] current ray queue, [2] next ray queue, [3] shadow queue ([2] and [3] are one 64-bit word for blockAppend2); stats: [0] camera rays, [1] closest-hit queries, [2] shadow queries
__global__ void k_advance(uint32_t *counters, unsigned long long *stats, int what) {
	if (what == 0) { stats[0] += counters[0]; return; }                     // after k_generate
	if (what == 1) {                                                        // path: one bounce done
		stats[1] += counters[0]; stats[2] += counters[3];
		counters[0] = counters[2]; counters[2] = 0; counters[3] = 0;
		return;
	}
	if (what == 2) { stats[1] += counters[0]; return; }                     // direct: primary hits traced
	stats[1] += counters[2]; stats[2] += counters[3];                       // direct: one round done (the primary queue stays)
	counters[2] = 0; counters[3] = 0;
}

// ImageBlock::putSample (imageblock.h:80-138) for all samples of one tile, accumulated in shared
// memory over the tile plus the filter border, then flushed once.
__global__ void __launch_bounds__(TILE_PIX)
k_splat(const FilmDev film, const BatchDesc b, const float4 *__restrict__ li, const float2 *__restrict__ samplePos,
		const float *__restrict__ filterTable, float *filmAcc) {
	extern __shared__ float acc[];          // (TILE + 2*border)^2 * 5
	__shared__ float table[256];
	const int border = film.border;
	const int side = TILE + 2 * border;
	const int nAcc = side * side * 5;
	for (int k = threadIdx.x; k < nAcc; k += TILE_PIX) acc[k] = 0.0f;
	table[threadIdx.x] = filterTable[threadIdx.x];
	__syncthreads();

	const uint32_t tile = b.tileBegin + blockIdx.x;
	const int tx0 = (int) (tile % b.tilesX) * TILE, ty0 = (int) (tile / b.tilesX) * TILE;   // crop-relative
	const size_t base = (size_t) blockIdx.x * b.sampleCount * TILE_PIX;
	const uint32_t nSlots = b.sampleCount * TILE_PIX;
	const bool box = (border == 0);
	for (uint32_t k = threadIdx.x; k < nSlots; k += TILE_PIX) {
		const float2 sp = samplePos[base + k];
		if (sp.x != sp.x) continue;                              // slot outside the crop window
		const float4 v = li[base + k];
		// !spec.isValid(): NaN or negative components drop the sample (imageblock.h:85-88)
		if (v.x != v.x || v.y != v.y || v.z != v.z || v.x < 0.0f || v.y < 0.0f || v.z < 0.0f) continue;
		// position relative to this tile's accumulator (block offset = tile origin, same border)
		const float fx = sp.x - 0.5f - (float) (film.cropX + tx0 - border);
		const float fy = sp.y - 0.5f - (float) (film.cropY + ty0 - border);
		int xStart = (int) ceilf(fx - film.filterSizeX), xEnd = (int) floorf(fx + film.filterSizeX);
		int yStart = (int) ceilf(fy - film.filterSizeY), yEnd = (int) floorf(fy + film.filterSizeY);
		xStart = max(0, xStart); yStart = max(0, yStart);
		xEnd = min(xEnd, side - 1); yEnd = min(yEnd, side - 1);
		for (int y = yStart; y <= yEnd; ++y) {
			const int iy = min((int) (film.factorY * fabsf((float) y - fy)), 15);
			for (int x = xStart; x <= xEnd; ++x) {
				const int ix = min((int) (film.factorX * fabsf((float) x - fx)), 15);
				const float w = table[iy * 16 + ix];
				float *a = &acc[(y * side + x) * 5];
				// shared-memory atomics: other warps splat other sample indices into the same pixels
				atomicAdd(a + 0, v.x * w); atomicAdd(a + 1, v.y * w); atomicAdd(a + 2, v.z * w);
				atomicAdd(a + 3, v.w * w); atomicAdd(a + 4, w);
			}
		}
	}
	__syncthreads();
	// flush: consecutive threads write consecutive floats of a row (5 floats per pixel)
	const int rowFloats = side * 5;
	for (int k = threadIdx.x; k < nAcc; k += TILE_PIX) {
		const int y = k / rowFloats, xf = k % rowFloats;
		const int gx = tx0 - border + xf / 5, gy = ty0 - border + y;    // crop-relative pixel
		if (gx < 0 || gy < 0 || gx >= film.cropW || gy >= film.cropH) continue;
		const float val = acc[k];
		float *dst = &filmAcc[((size_t) gy * film.cropW + gx) * 5 + xf % 5];
		if (box) *dst += val;            // tiles are disjoint under the box filter: plain coalesced RMW
		else if (val != 0.0f) atomicAdd(dst, val);   // borders overlap neighbouring tiles
	}
}

__global__ void k_zero(float *p, size_t n) {
	const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) p[i] = 0.0f;
}

inline unsigned blocksFor(size_t n, unsigned bs) { return (unsigned) ((n + bs - 1) / bs); }

} // namespace

// Launch helpers: pick the traversal kernel variant per queue. `nPtr`: the queue's length in device memory (the host does not
// know it); grids are one resident wave, the kernels loop or pull work until the queue is empty.
static unsigned fullGrid(mtsgpu_ctx *ctx, unsigned perSM) {
	(void) tracePersistentBlocks(ctx, 1);          // fills ctx->smCount
	return (unsigned) ctx->smCount * perSM;
}
static int launchClosestWf(mtsgpu_ctx *ctx, const float4 *rays, const uint32_t *nPtr, uint4 *hits, bool persistent) {
	if (ctx->scene.hasInstances) {     // instance hits also report the nested triangle: the lock-step kernel writes it behind the hits
		k_closest_wf<true><<<fullGrid(ctx, 16), TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, rays, nPtr, hits, reinterpret_cast<uint32_t *>(hits + ctx->pathCapacity));
		return MTSGPU_OK;
	}
	if (persistent) {
		uint32_t *rc = nullptr; int e = traceRayCounter(ctx, &rc); if (e) return e;
		const unsigned blocks = fullGrid(ctx, TRACE_PERSISTENT_BLOCKS_PER_SM);
		const int rb = traceRefetchBelow(ctx), lt = traceLeafThreshold(ctx);
		unsigned long long *cnt = ctx->countersEnabled ? ctx->dCounters.as<unsigned long long>() + 2 : nullptr;     // [2..3]: closest-hit queues of a render
#define WFP(S, C) k_closest_wfp<S, C><<<blocks, TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, rays, nPtr, hits, rc, rb, lt, cnt)
		if (ctx->scene.hasSpheres) { if (cnt) WFP(true, true); else WFP(true, false); }
		else { if (cnt) WFP(false, true); else WFP(false, false); }
#undef WFP
	} else {
		if (ctx->scene.hasSpheres) k_closest_wf<true><<<fullGrid(ctx, 16), TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, rays, nPtr, hits, nullptr);
		else k_closest_wf<false><<<fullGrid(ctx, 16), TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, rays, nPtr, hits, nullptr);
	}
	return MTSGPU_OK;
}
static int launchShadowWf(mtsgpu_ctx *ctx, const float4 *rays, const float4 *payload, const uint32_t *nPtr, float4 *li, bool persistent) {
	if (persistent) {
		uint32_t *rc = nullptr; int e = traceRayCounter(ctx, &rc); if (e) return e;
		const unsigned blocks = fullGrid(ctx, TRACE_PERSISTENT_BLOCKS_PER_SM);
		const int rb = traceRefetchBelow(ctx), lt = traceLeafThreshold(ctx);
		unsigned long long *cnt = ctx->countersEnabled ? ctx->dCounters.as<unsigned long long>() + 4 : nullptr;     // [4..5]: shadow queues of a render
#define WFP(S, C) k_shadow_wfp<S, C><<<blocks, TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, rays, payload, nPtr, li, rc, rb, lt, cnt)
		if (ctx->scene.hasSpheres) { if (cnt) WFP(true, true); else WFP(true, false); }
		else { if (cnt) WFP(false, true); else WFP(false, false); }
#undef WFP
	} else {
		if (ctx->scene.hasSpheres) k_shadow_wf<true><<<fullGrid(ctx, 16), TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, rays, payload, nPtr, li);
		else k_shadow_wf<false><<<fullGrid(ctx, 16), TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, rays, payload, nPtr, li);
	}
	return MTSGPU_OK;
}

static double wallSeconds() {
	return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static int ensureWavefront(mtsgpu_ctx *ctx, uint32_t capacity) {
	if (ctx->pathCapacity >= capacity) return MTSGPU_OK;
	const size_t c = capacity;
	CUDA_TRY(ctx, ctx->dPathState.alloc(c * (16 + 8)));            // per slot: li + sample position
	CUDA_TRY(ctx, ctx->dRayA.alloc(c * 32)); CUDA_TRY(ctx, ctx->dRayB.alloc(c * 32));
	CUDA_TRY(ctx, ctx->dStateA.alloc(c * 32)); CUDA_TRY(ctx, ctx->dStateB.alloc(c * 32));     // per queue position, beside the rays
	CUDA_TRY(ctx, ctx->dHit.alloc(c * 20));                 // uint4 hits + one nested-triangle index each
	CUDA_TRY(ctx, ctx->dShadowRay.alloc(c * 32)); CUDA_TRY(ctx, ctx->dShadowPayload.alloc(c * 16));
	CUDA_TRY(ctx, ctx->dQueueA.alloc(c * 4)); CUDA_TRY(ctx, ctx->dQueueB.alloc(c * 4));
	CUDA_TRY(ctx, ctx->dQueueCounts.alloc(64));
	ctx->pathCapacity = capacity;
	return MTSGPU_OK;
}

static int uploadFilter(mtsgpu_ctx *ctx, DevBuf &dTable) {
	CUDA_TRY(ctx, dTable.alloc(256 * sizeof(float)));
	CUDA_TRY(ctx, cudaMemcpyAsync(dTable.p, ctx->filterTable, 256 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	return MTSGPU_OK;
}

static int makeSamplerConfig(mtsgpu_ctx *ctx, const mtsgpu_render_params *P, SamplerConfig &cfg) {
	REQUIRE(ctx, P->samplerType <= MTSGPU_SAMPLER_HAMMERSLEY, MTSGPU_ERR_UNSUPPORTED, "render: unknown sampler type %u", P->samplerType);
	if ((P->samplerType == MTSGPU_SAMPLER_HALTON || P->samplerType == MTSGPU_SAMPLER_HAMMERSLEY) && P->mode == MTSGPU_MODE_DIRECT)      // halton.cpp:99-105, hammersley.cpp:111-117
		REQUIRE(ctx, P->luminaireSamples == 1 && P->bsdfSamples == 1, MTSGPU_ERR_UNSUPPORTED, "request2DArray() is not supported by QMC samplers!");
	cfg.type = P->samplerType; cfg.depth = P->samplerDepth; cfg.log2Count = 0; cfg.resolution = 1;
	cfg.lensDraw = (ctx->haveCamera && ctx->camera.apertureRadius > 0.0f) ? 1u : 0u;
	cfg.timeDraw = (ctx->haveCamera && ctx->camera.shutterOpenTime > 0.0f) ? 1u : 0u;      // Camera::needsTimeSample (camera.h:64)
	if (cfg.type == MTSGPU_SAMPLER_HAMMERSLEY) {
		REQUIRE(ctx, P->sppTotal > 0, MTSGPU_ERR_ARG, "render: hammersley needs the sample count");
		cfg.resolution = P->sppTotal;                  // m_invSamplesPerPixel = 1 / sampleCount (hammersley.cpp:46-47)
	}
	if (cfg.type == MTSGPU_SAMPLER_LD) {
		// the reference rounds the count up to a power of two and renders that many samples (ldsampler.cpp:61-65)
		REQUIRE(ctx, P->sppTotal > 0 && (P->sppTotal & (P->sppTotal - 1)) == 0, MTSGPU_ERR_ARG,
			"render: Sample count should be a power of two (got %u)", P->sppTotal);
		while ((1u << cfg.log2Count) < P->sppTotal) cfg.log2Count++;
	}
	if (cfg.type == MTSGPU_SAMPLER_STRATIFIED) {
		// the reference rounds the count up to a perfect square and renders that many samples (stratified.cpp:37-47)
		uint32_t r = 1;
		while ((uint64_t) r * r < P->sppTotal) ++r;
		REQUIRE(ctx, P->sppTotal > 0 && r * r == P->sppTotal, MTSGPU_ERR_ARG, "render: Sample count should be a perfect square (got %u)", P->sppTotal);
		cfg.resolution = r;
		while ((1u << cfg.log2Count) < P->sppTotal) cfg.log2Count++;
	}
	return MTSGPU_OK;
}

int renderDevice(mtsgpu_ctx *ctx, const mtsgpu_render_params *P, volatile const int *cancel, float *dOutFilm, mtsgpu_stats *stats) {
	REQUIRE(ctx, ctx->built, MTSGPU_ERR_ARG, "render: call mtsgpu_build_accel first");
	REQUIRE(ctx, ctx->haveCamera && ctx->haveFilm, MTSGPU_ERR_ARG, "render: camera and film must be set");
	REQUIRE(ctx, P->mode == MTSGPU_MODE_PATH || P->mode == MTSGPU_MODE_DIRECT, MTSGPU_ERR_ARG, "render: unknown mode %d", P->mode);
	REQUIRE(ctx, P->sppBegin <= P->sppEnd && P->sppEnd <= P->sppTotal, MTSGPU_ERR_ARG, "render: bad sample range [%u, %u) of %u", P->sppBegin, P->sppEnd, P->sppTotal);
	if (P->mode == MTSGPU_MODE_DIRECT)
		REQUIRE(ctx, P->luminaireSamples >= 0 && P->bsdfSamples >= 0 && P->luminaireSamples + P->bsdfSamples > 0, MTSGPU_ERR_ARG, "render: luminaireSamples + bsdfSamples must be > 0");
	SamplerConfig samplerConfig;
	{ int rc = makeSamplerConfig(ctx, P, samplerConfig); if (rc) return rc; }
	cudaStream_t st = ctx->stream;
	static const bool persistClosest = getenv("MTSGPU_WF_PERSIST_CLOSEST") ? atoi(getenv("MTSGPU_WF_PERSIST_CLOSEST")) != 0 : true;
	static const bool persistShadow = getenv("MTSGPU_WF_PERSIST_SHADOW") ? atoi(getenv("MTSGPU_WF_PERSIST_SHADOW")) != 0 : true;
	const FilmDev film = ctx->film;
	const size_t filmFloats = (size_t) film.cropW * film.cropH * 5;
	{
		int rc = rebuildLuminaireCdf(ctx, P->importanceSampleLuminaires != 0);
		if (rc) return rc;
	}
	DevBuf &dTable = ctx->dFilterTable;
	{ int rc = uploadFilter(ctx, dTable); if (rc) return rc; }

	if (ctx->countersEnabled) {
		CUDA_TRY(ctx, ctx->dCounters.alloc(64));
		CUDA_TRY(ctx, cudaMemsetAsync(ctx->dCounters.p, 0, 64, st));
	}
	k_zero<<<blocksFor(filmFloats, 256), 256, 0, st>>>(dOutFilm, filmFloats);
	CUDA_TRY(ctx, cudaGetLastError());
	mtsgpu_stats S; memset(&S, 0, sizeof(S));
	S.kernelLaunches = 1;

	const uint32_t tilesX = (film.cropW + TILE - 1) / TILE, tilesY = (film.cropH + TILE - 1) / TILE;
	const uint32_t nTiles = tilesX * tilesY;
	const uint32_t spp = P->sppEnd - P->sppBegin;
	cudaEvent_t evBegin = ctx->ev[0], evEnd = ctx->ev[1], evA = ctx->ev[2], evB = ctx->ev[3];
	CUDA_TRY(ctx, cudaEventRecord(evBegin, st));
	if (spp > 0) {
		const uint32_t slotsLog2 = getenv("MTSGPU_BATCH_LOG2") ? (uint32_t) std::min(std::max(atoi(getenv("MTSGPU_BATCH_LOG2")), 16), 28) : 27u;
		// 228 B of queues + path state per slot => 30.6 GB at the default size (2^27 slots), allocated only as far as the job needs.
		// Measured on C4 (B200): 2^25 -> 2^27 slots per batch = -11 % render time: every bounce of a batch is one launch of each
		// kernel, and the late bounces (1-2 % of the rays) and each launch's drain phase weigh less in a larger batch.
		const uint32_t maxSlots = 1u << slotsLog2;
		uint32_t sampleChunk = std::min<uint32_t>(spp, maxSlots / TILE_PIX);
		uint32_t tilesPerBatch = std::max<uint32_t>(1u, std::min<uint32_t>(nTiles, maxSlots / (TILE_PIX * sampleChunk)));
		const uint32_t capacity = tilesPerBatch * sampleChunk * TILE_PIX;
		{
			// size the queues for the whole job (sppTotal), so that a short warm-up call and the full render
			// share one allocation
			const uint64_t wholeJob = (uint64_t) nTiles * TILE_PIX * std::max<uint32_t>(P->sppTotal, spp);
			const uint32_t allocSlots = (uint32_t) std::min<uint64_t>(maxSlots, std::max<uint64_t>(capacity, wholeJob));
			int rc = ensureWavefront(ctx, allocSlots); if (rc) return rc;
		}
		DevBuf &dPrimSlot = ctx->dPrimSlot, &dPrimRay = ctx->dPrimRay, &dPrimHit = ctx->dPrimHit;
		if (P->mode == MTSGPU_MODE_DIRECT) {
			CUDA_TRY(ctx, dPrimSlot.alloc((size_t) ctx->pathCapacity * 4));
			CUDA_TRY(ctx, dPrimRay.alloc((size_t) ctx->pathCapacity * 32));
			CUDA_TRY(ctx, ctx->dPrimState.alloc((size_t) ctx->pathCapacity * 32));
			CUDA_TRY(ctx, dPrimHit.alloc((size_t) ctx->pathCapacity * 20));
		}
		// the camera behind DeviceScene::camera (ray differentials for bitmap textures)
		CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dCamera.p, &ctx->camera, sizeof(CameraDev), cudaMemcpyHostToDevice, st));
		WfArrays wf;
		char *ps = ctx->dPathState.as<char>();
		wf.li = (float4 *) ps;
		wf.samplePos = (float2 *) (ps + (size_t) capacity * 16);
		wf.statePrim = ctx->dPrimState.as<float4>();
		float4 *stateBuf[2] = { ctx->dStateA.as<float4>(), ctx->dStateB.as<float4>() };
		wf.hit = ctx->dHit.as<uint4>(); wf.capacity = ctx->pathCapacity;
		wf.shadowRay = ctx->dShadowRay.as<float4>(); wf.shadowPayload = ctx->dShadowPayload.as<float4>();
		wf.counters = ctx->dQueueCounts.as<uint32_t>();
		wf.slotPrim = dPrimSlot.as<uint32_t>(); wf.rayPrim = dPrimRay.as<float4>(); wf.hitPrim = dPrimHit.as<uint4>();
		uint32_t *slotBuf[2] = { ctx->dQueueA.as<uint32_t>(), ctx->dQueueB.as<uint32_t>() };
		float4 *rayBuf[2] = { ctx->dRayA.as<float4>(), ctx->dRayB.as<float4>() };

		bool bmp = false;
		for (const mtsgpu_texture &t : ctx->textures) bmp |= t.type == MTSGPU_TEXTURE_BITMAP;
		bmp |= ctx->scene.hasInstances != 0;          // the same variant also shades instance hits
		IntegratorDev ig;
		ig.mode = P->mode; ig.maxDepth = P->maxDepth; ig.rrDepth = P->rrDepth; ig.strictNormals = P->strictNormals;
		ig.nLum = P->luminaireSamples; ig.nBsdf = P->bsdfSamples; ig.importance = (int) P->importanceSampleLuminaires;
		if (P->mode == MTSGPU_MODE_DIRECT) ig.strictNormals = 0;     // MIDirectIntegrator has no such property
		const int splatSmem = (TILE + 2 * film.border) * (TILE + 2 * film.border) * 5 * (int) sizeof(float);

		// ---- no host round trip inside a batch: queue lengths stay in device memory (counters[0] current ray queue, [2] next ray
		// queue, [3] shadow queue; k_advance rotates them and keeps the statistics), every kernel reads its own length, and
		// the host only learns — a few bounces late, through an asynchronous copy it polls — when a batch has run dry.
		if (!ctx->hCounts) CUDA_TRY(ctx, cudaMallocHost(&ctx->hCounts, WF_POLL_RING * sizeof(uint32_t)));     // pinned, kept for the context's lifetime
		volatile uint32_t *hRing = ctx->hCounts;
		for (int i = 0; i < WF_POLL_RING; ++i)
			if (!ctx->pollEv[i]) CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->pollEv[i], cudaEventDisableTiming));
		CUDA_TRY(ctx, ctx->dWfStats.alloc(4 * sizeof(unsigned long long)));
		CUDA_TRY(ctx, cudaMemsetAsync(ctx->dWfStats.p, 0, 4 * sizeof(unsigned long long), st));
		unsigned long long *dStats = ctx->dWfStats.as<unsigned long long>();
		int rc = MTSGPU_OK;
		auto fail = [&](int code) { return code; };
		float msClosest = 0, msShadow = 0;
		// kernel-class timing: event pairs around the traversal launches, two sets used by alternate batches and read back
		// one batch late (by then they have long completed)
		std::vector<int> evKind[2]; size_t evUsed[2] = { 0, 0 };
		auto stamp = [&](int set, int kind) -> cudaError_t {        // kind: 0 / 1 = before / after closest-hit, 2 / 3 = shadow
			std::vector<cudaEvent_t> &pool = ctx->evPool[set];
			if (evUsed[set] == pool.size()) { cudaEvent_t e; cudaError_t er = cudaEventCreate(&e); if (er != cudaSuccess) return er; pool.push_back(e); }
			if (evKind[set].size() <= evUsed[set]) evKind[set].resize(evUsed[set] + 1);
			evKind[set][evUsed[set]] = kind;
			return cudaEventRecord(pool[evUsed[set]++], st);
		};
		auto harvest = [&](int set) -> cudaError_t {
			if (evUsed[set] == 0) return cudaSuccess;
			std::vector<cudaEvent_t> &pool = ctx->evPool[set];
			cudaError_t er = cudaEventSynchronize(pool[evUsed[set] - 1]);
			if (er != cudaSuccess) return er;
			for (size_t i = 0; i + 1 < evUsed[set]; i += 2) {
				float ms = 0; cudaEventElapsedTime(&ms, pool[i], pool[i + 1]);
				if (evKind[set][i] == 0) msClosest += ms; else msShadow += ms;
			}
			evUsed[set] = 0;
			return cudaSuccess;
		};
		const unsigned shadeGrid = fullGrid(ctx, SHADE_MIN_BLOCKS);
#define LAUNCH_SHADE(M, NIN, ROUND) do { if (bmp) k_shade<M, true><<<shadeGrid, WF_BLOCK, 0, st>>>(ctx->scene, ig, b, film, wf, (NIN), (ROUND)); \
			else k_shade<M, false><<<shadeGrid, WF_BLOCK, 0, st>>>(ctx->scene, ig, b, film, wf, (NIN), (ROUND)); S.kernelLaunches++; } while (0)
		uint32_t batchIndex = 0;
		// ---- progress / preview (mtsgpu_set_progress): a device-side snapshot of the film after a batch's splat, copied to pinned
		// memory on a stream of its own; the callback runs from this loop once the copy has landed (never waited for)
		const uint32_t batchesTotal = ((nTiles + tilesPerBatch - 1) / tilesPerBatch) * ((spp + sampleChunk - 1) / sampleChunk);
		bool progressPending = false; double progressFraction = 0, progressLast = wallSeconds();
		if (ctx->progressFn && !ctx->progressStream) {
			CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->progressStream, cudaStreamNonBlocking));
			CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->progressSnap, cudaEventDisableTiming));
			CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->progressCopied, cudaEventDisableTiming));
		}
		if (ctx->progressFn && ctx->progressWantFilm) {
			CUDA_TRY(ctx, ctx->dProgressSnap.alloc(filmFloats * sizeof(float)));
			if (ctx->hProgressBytes < filmFloats * sizeof(float)) {
				if (ctx->hProgressFilm) cudaFreeHost(ctx->hProgressFilm);
				ctx->hProgressFilm = nullptr; ctx->hProgressBytes = 0;
				CUDA_TRY(ctx, cudaMallocHost(&ctx->hProgressFilm, filmFloats * sizeof(float)));
				ctx->hProgressBytes = filmFloats * sizeof(float);
			}
		}
		auto progressDeliver = [&](bool wait) {
			if (!progressPending) return;
			if (wait) cudaEventSynchronize(ctx->progressCopied);
			else if (cudaEventQuery(ctx->progressCopied) != cudaSuccess) return;
			progressPending = false; progressLast = wallSeconds();
			ctx->progressFn(ctx->progressUser, progressFraction, ctx->progressWantFilm ? ctx->hProgressFilm : nullptr);
		};

		for (uint32_t t0 = 0; t0 < nTiles && rc == MTSGPU_OK; t0 += tilesPerBatch) {
			for (uint32_t s0 = 0; s0 < spp; s0 += sampleChunk, ++batchIndex) {
				if (cancel && *cancel) { rc = setError(ctx, MTSGPU_ERR_CANCELLED, "render cancelled"); break; }
				const int evSet = (int) (batchIndex & 1u);
				CUDA_TRY(ctx, harvest(evSet));                   // the events of the batch before the previous one
				BatchDesc b;
				b.tileBegin = t0; b.tileCount = std::min(tilesPerBatch, nTiles - t0); b.tilesX = tilesX;
				b.sampleBegin = P->sppBegin + s0; b.sampleCount = std::min(sampleChunk, spp - s0);
				b.slots = b.tileCount * b.sampleCount * TILE_PIX; b.seed = P->seed;
				b.sampler = samplerConfig;
				b.diffScale = 1.0f / std::sqrt((float) P->sppTotal);          // integrator.cpp:145
				int cur = 0;
				const bool direct = P->mode == MTSGPU_MODE_DIRECT;
				wf.slotNext = direct ? wf.slotPrim : slotBuf[cur]; wf.rayNext = direct ? wf.rayPrim : rayBuf[cur];
				wf.stateNext = direct ? wf.statePrim : stateBuf[cur]; wf.stateCur = stateBuf[cur];
				CUDA_TRY(ctx, cudaMemsetAsync(wf.counters, 0, 16, st));
				k_generate<<<blocksFor(b.slots, WF_BLOCK), WF_BLOCK, 0, st>>>(ctx->camera, film, b, wf);
				k_advance<<<1, 1, 0, st>>>(wf.counters, dStats, 0);
				S.kernelLaunches += 2;

				if (!direct) {
					// path.cpp:60-216: at most maxDepth closest-hit queries per path (the camera ray included); unbounded depth
					// runs until the polled queue length reads zero
					const int maxBounces = P->maxDepth > 0 ? P->maxDepth : (P->maxDepth == 0 ? 1 : -1);
					uint32_t enq = 0, seen = 0; bool dry = false;
					for (int bounce = 0; maxBounces < 0 || bounce < maxBounces; ++bounce) {
						if (cancel && *cancel) { rc = setError(ctx, MTSGPU_ERR_CANCELLED, "render cancelled"); break; }
						while (seen < enq) {                 // finished polls: never wait unless the host is WF_POLL_AHEAD bounces ahead
							cudaEvent_t pe = ctx->pollEv[seen % WF_POLL_RING];
							if (enq - seen >= WF_POLL_AHEAD) CUDA_TRY(ctx, cudaEventSynchronize(pe));
							else if (cudaEventQuery(pe) != cudaSuccess) break;
							if (hRing[seen % WF_POLL_RING] == 0) dry = true;
							seen++;
						}
						if (dry) break;
						progressDeliver(false);
						wf.slotCur = slotBuf[cur]; wf.rayCur = rayBuf[cur]; wf.stateCur = stateBuf[cur];
						wf.slotNext = slotBuf[cur ^ 1]; wf.rayNext = rayBuf[cur ^ 1]; wf.stateNext = stateBuf[cur ^ 1];
						CUDA_TRY(ctx, stamp(evSet, 0));
						if (launchClosestWf(ctx, wf.rayCur, wf.counters, wf.hit, persistClosest)) return fail(MTSGPU_ERR_CUDA);
						CUDA_TRY(ctx, stamp(evSet, 1));
						LAUNCH_SHADE(0, wf.counters, 0u);
						CUDA_TRY(ctx, stamp(evSet, 2));
						if (launchShadowWf(ctx, wf.shadowRay, wf.shadowPayload, wf.counters + 3, wf.li, persistShadow)) return fail(MTSGPU_ERR_CUDA);
						CUDA_TRY(ctx, stamp(evSet, 3));
						k_advance<<<1, 1, 0, st>>>(wf.counters, dStats, 1);
						S.kernelLaunches += 3;
						CUDA_TRY(ctx, cudaMemcpyAsync((void *) (hRing + enq % WF_POLL_RING), wf.counters, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
						CUDA_TRY(ctx, cudaEventRecord(ctx->pollEv[enq % WF_POLL_RING], st));
						enq++;
						cur ^= 1;
					}
				} else {
					// primary hit once, kept in the *Prim buffers (k_generate wrote them); then `rounds` passes of (NEE, BSDF ray)
					CUDA_TRY(ctx, stamp(evSet, 0));
					if (launchClosestWf(ctx, wf.rayPrim, wf.counters, wf.hitPrim, persistClosest)) return fail(MTSGPU_ERR_CUDA);
					CUDA_TRY(ctx, stamp(evSet, 1));
					k_advance<<<1, 1, 0, st>>>(wf.counters, dStats, 2);
					S.kernelLaunches += 2;
					const int rounds = std::max(P->luminaireSamples, P->bsdfSamples);
					for (int r = 0; r < rounds; ++r) {
						if (cancel && *cancel) { rc = setError(ctx, MTSGPU_ERR_CANCELLED, "render cancelled"); break; }
						wf.slotNext = slotBuf[0]; wf.rayNext = rayBuf[0]; wf.stateNext = stateBuf[0];
						wf.slotCur = slotBuf[0]; wf.rayCur = rayBuf[0]; wf.stateCur = stateBuf[0];
						LAUNCH_SHADE(1, wf.counters, (uint32_t) r);
						CUDA_TRY(ctx, stamp(evSet, 2));
						if (launchShadowWf(ctx, wf.shadowRay, wf.shadowPayload, wf.counters + 3, wf.li, persistShadow)) return fail(MTSGPU_ERR_CUDA);
						CUDA_TRY(ctx, stamp(evSet, 3));
						CUDA_TRY(ctx, stamp(evSet, 0));
						if (launchClosestWf(ctx, wf.rayCur, wf.counters + 2, wf.hit, persistClosest)) return fail(MTSGPU_ERR_CUDA);
						CUDA_TRY(ctx, stamp(evSet, 1));
						LAUNCH_SHADE(2, wf.counters + 2, (uint32_t) r);
						k_advance<<<1, 1, 0, st>>>(wf.counters, dStats, 3);
						S.kernelLaunches += 3;
					}
				}
				if (rc != MTSGPU_OK) break;
				k_splat<<<b.tileCount, TILE_PIX, splatSmem, st>>>(film, b, wf.li, wf.samplePos, dTable.as<float>(), dOutFilm);
				S.kernelLaunches++;
				CUDA_TRY(ctx, cudaGetLastError());
				if (ctx->progressFn) {
					progressDeliver(false);
					if (!progressPending && batchIndex + 1 < batchesTotal && wallSeconds() - progressLast >= ctx->progressInterval) {
						progressFraction = (double) (batchIndex + 1) / (double) batchesTotal;
						if (ctx->progressWantFilm) {
							CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dProgressSnap.p, dOutFilm, filmFloats * sizeof(float), cudaMemcpyDeviceToDevice, st));
							CUDA_TRY(ctx, cudaEventRecord(ctx->progressSnap, st));
							CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->progressStream, ctx->progressSnap, 0));
							CUDA_TRY(ctx, cudaMemcpyAsync(ctx->hProgressFilm, ctx->dProgressSnap.p, filmFloats * sizeof(float), cudaMemcpyDeviceToHost, ctx->progressStream));
							CUDA_TRY(ctx, cudaEventRecord(ctx->progressCopied, ctx->progressStream));
						} else {
							CUDA_TRY(ctx, cudaEventRecord(ctx->progressCopied, st));
						}
						progressPending = true;
					}
				}
			}
		}
		if (ctx->progressFn) progressDeliver(true);
#undef LAUNCH_SHADE
		CUDA_TRY(ctx, harvest(0)); CUDA_TRY(ctx, harvest(1));
		{
			unsigned long long h[4];
			CUDA_TRY(ctx, cudaMemcpyAsync(h, dStats, sizeof(h), cudaMemcpyDeviceToHost, st));
			CUDA_TRY(ctx, cudaStreamSynchronize(st));
			S.cameraRays = h[0]; S.normalRays = h[1]; S.shadowRays = h[2];
		}
		S.samples = S.cameraRays;
		S.traceClosestMs = msClosest; S.traceAnyMs = msShadow;
		if (rc != MTSGPU_OK) return fail(rc);
	}
	CUDA_TRY(ctx, cudaEventRecord(evEnd, st));
	CUDA_TRY(ctx, cudaStreamSynchronize(st));
	CUDA_TRY(ctx, cudaGetLastError());
	{ float ms = 0; cudaEventElapsedTime(&ms, evBegin, evEnd); S.totalMs = ms; }
	S.shadeMs = S.totalMs - S.traceClosestMs - S.traceAnyMs;
	if (ctx->countersEnabled) {
		unsigned long long h[8];
		CUDA_TRY(ctx, cudaMemcpy(h, ctx->dCounters.p, sizeof(h), cudaMemcpyDeviceToHost));
		S.closestNodeVisits = h[2]; S.closestTriTests = h[3]; S.shadowNodeVisits = h[4]; S.shadowTriTests = h[5];
	}
	if (stats) *stats = S;
	{ const int frc = checkTraceFlags(ctx); if (frc) return frc; }
	if (ctx->progressFn) ctx->progressFn(ctx->progressUser, 1.0, nullptr);
	return MTSGPU_OK;
}

// ================================================================================================
// Per-function test hooks (C ABI: mtsgpu_bsdf_eval, mtsgpu_emitter_sample, mtsgpu_camera_rays,
// mtsgpu_film_splat, mtsgpu_intersect_full). They run the same device functions as the render loop.
// ================================================================================================
namespace {

__global__ void k_hook_bsdf(const DeviceScene sc, const float *q, uint32_t n, uint32_t nMat, float *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float *e = q + (size_t) i * 9;
	float *o = out + (size_t) i * 16;
	for (int k = 0; k < 16; ++k) o[k] = 0;
	const uint32_t id = (uint32_t) e[0];
	if (id >= nMat) return;
	const MaterialDev m = sc.materials[id];
	const V3 wi = v3(e[1], e[2], e[3]), wo = v3(e[4], e[5], e[6]);
	const V3 f = bsdfF(m, wi, wo);
	o[0] = f.x; o[1] = f.y; o[2] = f.z; o[3] = bsdfPdf(m, wi, wo);
	V3 swo = v3(0, 0, 0); float pdf = 0; uint32_t type = 0;
	const V3 s3 = bsdfSample3(m, wi, e[7], e[8], swo, pdf, type);
	if (!isZero(s3)) {
		o[4] = swo.x; o[5] = swo.y; o[6] = swo.z; o[7] = s3.x; o[8] = s3.y; o[9] = s3.z; o[10] = pdf; o[11] = (float) type;
	}
	const V3 s2 = bsdfSample2(m, wi, e[7], e[8], swo, type);
	o[12] = s2.x; o[13] = s2.y; o[14] = s2.z;
}

__global__ void k_hook_emitter(const DeviceScene sc, const float *q, uint32_t n, int importance, float *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float *e = q + (size_t) i * 5;
	float *o = out + (size_t) i * 16;
	for (int k = 0; k < 16; ++k) o[k] = 0;
	const V3 p = v3(e[0], e[1], e[2]);
	LumSample ls;
	ls.p = ls.n = ls.d = ls.value = v3(0, 0, 0);
	sampleLuminaire(sc, p, e[3], e[4], importance != 0, ls);
	o[0] = ls.p.x; o[1] = ls.p.y; o[2] = ls.p.z; o[3] = ls.n.x; o[4] = ls.n.y; o[5] = ls.n.z;
	if (ls.pdf != 0) {
		o[6] = ls.d.x; o[7] = ls.d.y; o[8] = ls.d.z; o[9] = ls.pdf;
		o[10] = ls.value.x; o[11] = ls.value.y; o[12] = ls.value.z; o[13] = (float) ls.emitter;
		// Scene::pdfLuminaire (scene.cpp:381-394): zero for the delta luminaires
		const uint32_t type = sc.emitters[ls.emitter].type;
		o[14] = type == MTSGPU_EMITTER_AREA ? pdfLuminaire(sc, p, ls.p, ls.n, ls.emitter, importance != 0)
			: ((type == MTSGPU_EMITTER_CONSTANT || type == MTSGPU_EMITTER_ENVMAP) ? pdfBackground(sc, importance != 0, -ls.d) : 0.0f);
	} else {
		o[13] = -1.0f;
	}
}

__global__ void k_hook_camera(const CameraDev cam, const float *q, uint32_t n, float *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	V3 o, d; float mint, maxt;
	cameraRay(cam, q[4 * (size_t) i], q[4 * (size_t) i + 1], q[4 * (size_t) i + 2], q[4 * (size_t) i + 3], o, d, mint, maxt);
	float *r = out + (size_t) i * 8;
	r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = mint; r[4] = d.x; r[5] = d.y; r[6] = d.z; r[7] = maxt;
}

__global__ void k_hook_uv(const DeviceScene sc, const float4 *rays, const mtsgpu_hit *hits, uint32_t n, float *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float *o = out + (size_t) i * 8;
	for (int k = 0; k < 8; ++k) o[k] = 0;
	const mtsgpu_hit h = hits[i];
	if (h.shape == 0xFFFFFFFFu) return;
	Isect its;
	fillIntersection(sc, h.shape, h.prim, h.u, h.v, v3(rays[2 * (size_t) i]), v3(rays[2 * (size_t) i + 1]), h.t, its);
	o[0] = its.uvx; o[1] = its.uvy;
	if (its.materialId < 0) return;
	MaterialDev m = sc.materials[its.materialId];
	resolveTextures(sc, m, its.uvx, its.uvy);
	for (int k = 0; k < 6; ++k) o[2 + k] = m.p[k];
}

// camera samples [x, y, lens.x, lens.y] -> rays
__global__ void k_hook_camera_rays(const CameraDev cam, const float *q, uint32_t n, float4 *rays) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	V3 o, d; float mint, maxt;
	cameraRay(cam, q[4 * (size_t) i], q[4 * (size_t) i + 1], q[4 * (size_t) i + 2], q[4 * (size_t) i + 3], o, d, mint, maxt);
	rays[2 * (size_t) i] = make_float4(o.x, o.y, o.z, mint); rays[2 * (size_t) i + 1] = make_float4(d.x, d.y, d.z, maxt);
}

// rows of 12: its.uv, dudx, dudy, dvdx, dvdy, the two spectrum slots of the hit's BSDF with their textures looked up
__global__ void k_hook_texture(const DeviceScene sc, const float *q, const float4 *rays, const mtsgpu_hit *hits, uint32_t n, float diffScale, float *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float *o = out + (size_t) i * 12;
	for (int k = 0; k < 12; ++k) o[k] = 0;
	const mtsgpu_hit h = hits[i];
	if (h.shape == 0xFFFFFFFFu) return;
	Isect its;
	fillIntersection(sc, h.shape, h.prim, h.u, h.v, v3(rays[2 * (size_t) i]), v3(rays[2 * (size_t) i + 1]), h.t, its);
	o[0] = its.uvx; o[1] = its.uvy;
	if (its.materialId < 0) return;
	MaterialDev m = sc.materials[its.materialId];
	float4 duv = make_float4(0, 0, 0, 0);
	if (usesRayDifferentials(sc, m)) {
		const ShapeInfo si = sc.shapes[h.shape];
		const uint32_t tri = __ldg(sc.sortedOf + si.triOffset + (si.isSphere ? 0u : h.prim));
		duv = cameraHitPartials(sc, q[4 * (size_t) i], q[4 * (size_t) i + 1], q[4 * (size_t) i + 2], q[4 * (size_t) i + 3],
			v3(rays[2 * (size_t) i]), v3(rays[2 * (size_t) i + 1]), diffScale, tri, h.u, h.v, its.p, its.geoN);
	}
	o[2] = duv.x; o[3] = duv.y; o[4] = duv.z; o[5] = duv.w;
	resolveTextures(sc, m, its.uvx, its.uvy, duv);
	for (int k = 0; k < 6; ++k) o[6 + k] = m.p[k];
}

__global__ void k_hook_isect(const DeviceScene sc, const float4 *rays, const mtsgpu_hit *hits, uint32_t n, float *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float *o = out + (size_t) i * 12;
	const mtsgpu_hit h = hits[i];
	if (h.shape == 0xFFFFFFFFu) { for (int k = 0; k < 12; ++k) o[k] = 0; return; }
	const float4 a = rays[2 * (size_t) i], b = rays[2 * (size_t) i + 1];
	Isect its;
	fillIntersection(sc, h.shape, h.prim, h.u, h.v, v3(a), v3(b), h.t, its);
	o[0] = its.p.x; o[1] = its.p.y; o[2] = its.p.z; o[3] = its.geoN.x; o[4] = its.geoN.y; o[5] = its.geoN.z;
	o[6] = its.shN.x; o[7] = its.shN.y; o[8] = its.shN.z; o[9] = its.wi.x; o[10] = its.wi.y; o[11] = its.wi.z;
}

} // namespace

static int upload(mtsgpu_ctx *ctx, DevBuf &buf, const void *src, size_t bytes) {
	CUDA_TRY(ctx, buf.alloc(bytes));
	if (bytes) CUDA_TRY(ctx, cudaMemcpyAsync(buf.p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
	return MTSGPU_OK;
}
static int download(mtsgpu_ctx *ctx, void *dst, const DevBuf &buf, size_t bytes) {
	if (bytes) CUDA_TRY(ctx, cudaMemcpyAsync(dst, buf.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	CUDA_TRY(ctx, cudaGetLastError());
	return checkTraceFlags(ctx);
}

namespace {
__global__ void k_hook_sampler(SamplerConfig cfg, uint32_t pixel, uint32_t draw, uint32_t seed, uint32_t n, float2 *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	Sampler smp; smp.init(pixel, i, seed, cfg);
	// dimension whose draw number is `draw`: lens, pixel, then the per-bounce luminaire / BSDF pairs
	uint32_t dim;
	if (cfg.lensDraw && draw == 0) dim = DIM_LENS;
	else if (draw == cfg.lensDraw) dim = DIM_PIXEL;
	else { const uint32_t k = draw - cfg.lensDraw - 1; dim = DIM_BOUNCE(k / 2, k % 2); }
	float x, y; smp.get2D(dim, x, y);
	out[i] = make_float2(x, y);
}
}

int hookSamplerValues(mtsgpu_ctx *ctx, const mtsgpu_render_params *P, uint32_t pixel, uint32_t draw, float *out) {
	SamplerConfig cfg;
	{ int rc = makeSamplerConfig(ctx, P, cfg); if (rc) return rc; }
	const uint32_t n = P->sppTotal;
	if (n == 0) return MTSGPU_OK;
	CUDA_TRY(ctx, ctx->dStageOut.alloc((size_t) n * 8));
	k_hook_sampler<<<blocksFor(n, 128), 128, 0, ctx->stream>>>(cfg, pixel, draw, P->seed, n, ctx->dStageOut.as<float2>());
	return download(ctx, out, ctx->dStageOut, (size_t) n * 8);
}

int hookBsdfEval(mtsgpu_ctx *ctx, const float *queries, size_t n, float *out) {
	REQUIRE(ctx, ctx->built, MTSGPU_ERR_ARG, "bsdf_eval: call mtsgpu_build_accel first");
	if (n == 0) return MTSGPU_OK;
	int rc;
	if ((rc = upload(ctx, ctx->dStage, queries, n * 9 * 4))) return rc;
	CUDA_TRY(ctx, ctx->dStageOut.alloc(n * 16 * 4));
	k_hook_bsdf<<<blocksFor(n, 128), 128, 0, ctx->stream>>>(ctx->scene, ctx->dStage.as<float>(), (uint32_t) n,
		(uint32_t) ctx->materials.size(), ctx->dStageOut.as<float>());
	return download(ctx, out, ctx->dStageOut, n * 16 * 4);
}

int hookEmitterSample(mtsgpu_ctx *ctx, const float *queries, size_t n, float *out) {
	REQUIRE(ctx, ctx->built, MTSGPU_ERR_ARG, "emitter_sample: call mtsgpu_build_accel first");
	if (n == 0) return MTSGPU_OK;
	int rc;
	if ((rc = rebuildLuminaireCdf(ctx, true))) return rc;
	if ((rc = upload(ctx, ctx->dStage, queries, n * 5 * 4))) return rc;
	CUDA_TRY(ctx, ctx->dStageOut.alloc(n * 16 * 4));
	k_hook_emitter<<<blocksFor(n, 128), 128, 0, ctx->stream>>>(ctx->scene, ctx->dStage.as<float>(), (uint32_t) n, 1, ctx->dStageOut.as<float>());
	return download(ctx, out, ctx->dStageOut, n * 16 * 4);
}

int hookCameraRays(mtsgpu_ctx *ctx, const float *samples, size_t n, float *out) {
	REQUIRE(ctx, ctx->haveCamera, MTSGPU_ERR_ARG, "camera_rays: call mtsgpu_set_camera first");
	if (n == 0) return MTSGPU_OK;
	int rc;
	if ((rc = upload(ctx, ctx->dStage, samples, n * 4 * 4))) return rc;
	CUDA_TRY(ctx, ctx->dStageOut.alloc(n * 8 * 4));
	k_hook_camera<<<blocksFor(n, 128), 128, 0, ctx->stream>>>(ctx->camera, ctx->dStage.as<float>(), (uint32_t) n, ctx->dStageOut.as<float>());
	return download(ctx, out, ctx->dStageOut, n * 8 * 4);
}

int hookIntersectFull(mtsgpu_ctx *ctx, const float *rays, size_t n, mtsgpu_hit *hits, float *out) {
	REQUIRE(ctx, ctx->built, MTSGPU_ERR_ARG, "intersect_full: call mtsgpu_build_accel first");
	if (n == 0) return MTSGPU_OK;
	int rc;
	if ((rc = upload(ctx, ctx->dRays, rays, n * 32))) return rc;
	CUDA_TRY(ctx, ctx->dHits.alloc(n * sizeof(mtsgpu_hit)));
	if ((rc = launchTrace(ctx, ctx->dRays.as<float4>(), n, MTSGPU_TRACE_CLOSEST, ctx->dHits.p, nullptr))) return rc;
	CUDA_TRY(ctx, ctx->dStageOut.alloc(n * 12 * 4));
	k_hook_isect<<<blocksFor(n, 128), 128, 0, ctx->stream>>>(ctx->scene, ctx->dRays.as<float4>(), ctx->dHits.as<mtsgpu_hit>(), (uint32_t) n, ctx->dStageOut.as<float>());
	if (hits) CUDA_TRY(ctx, cudaMemcpyAsync(hits, ctx->dHits.p, n * sizeof(mtsgpu_hit), cudaMemcpyDeviceToHost, ctx->stream));
	return download(ctx, out, ctx->dStageOut, n * 12 * 4);
}

int hookIntersectUv(mtsgpu_ctx *ctx, const float *rays, size_t n, float *out) {
	REQUIRE(ctx, ctx->built, MTSGPU_ERR_ARG, "intersect_uv: call mtsgpu_build_accel first");
	if (n == 0) return MTSGPU_OK;
	int rc;
	if ((rc = upload(ctx, ctx->dRays, rays, n * 32))) return rc;
	CUDA_TRY(ctx, ctx->dHits.alloc(n * sizeof(mtsgpu_hit)));
	if ((rc = launchTrace(ctx, ctx->dRays.as<float4>(), n, MTSGPU_TRACE_CLOSEST, ctx->dHits.p, nullptr))) return rc;
	CUDA_TRY(ctx, ctx->dStageOut.alloc(n * 8 * 4));
	k_hook_uv<<<blocksFor(n, 128), 128, 0, ctx->stream>>>(ctx->scene, ctx->dRays.as<float4>(), ctx->dHits.as<mtsgpu_hit>(), (uint32_t) n, ctx->dStageOut.as<float>());
	return download(ctx, out, ctx->dStageOut, n * 8 * 4);
}

int hookTextureLookup(mtsgpu_ctx *ctx, const float *samples, size_t n, int sampleCount, float *out) {
	REQUIRE(ctx, ctx->built && ctx->haveCamera, MTSGPU_ERR_ARG, "texture_lookup: call mtsgpu_set_camera and mtsgpu_build_accel first");
	if (n == 0) return MTSGPU_OK;
	int rc;
	if ((rc = upload(ctx, ctx->dStage, samples, n * 16))) return rc;
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dCamera.p, &ctx->camera, sizeof(CameraDev), cudaMemcpyHostToDevice, ctx->stream));
	CUDA_TRY(ctx, ctx->dRays.alloc(n * 32));
	CUDA_TRY(ctx, ctx->dHits.alloc(n * sizeof(mtsgpu_hit)));
	k_hook_camera_rays<<<blocksFor(n, 128), 128, 0, ctx->stream>>>(ctx->camera, ctx->dStage.as<float>(), (uint32_t) n, ctx->dRays.as<float4>());
	if ((rc = launchTrace(ctx, ctx->dRays.as<float4>(), n, MTSGPU_TRACE_CLOSEST, ctx->dHits.p, nullptr))) return rc;
	CUDA_TRY(ctx, ctx->dStageOut.alloc(n * 12 * 4));
	k_hook_texture<<<blocksFor(n, 128), 128, 0, ctx->stream>>>(ctx->scene, ctx->dStage.as<float>(), ctx->dRays.as<float4>(), ctx->dHits.as<mtsgpu_hit>(), (uint32_t) n,
		1.0f / std::sqrt((float) std::max(sampleCount, 1)), ctx->dStageOut.as<float>());
	return download(ctx, out, ctx->dStageOut, n * 12 * 4);
}

// Film splat of an arbitrary sample list through the render loop's own k_splat: samples are
// bucketed by the tile that holds their pixel, padded to whole 256-slot passes.
int hookFilmSplat(mtsgpu_ctx *ctx, const float *samples, size_t n, float *outFilm) {
	REQUIRE(ctx, ctx->haveFilm, MTSGPU_ERR_ARG, "film_splat: call mtsgpu_set_film first");
	const FilmDev film = ctx->film;
	const uint32_t tilesX = (film.cropW + TILE - 1) / TILE, tilesY = (film.cropH + TILE - 1) / TILE;
	const uint32_t nTiles = tilesX * tilesY;
	std::vector<std::vector<uint32_t> > buckets(nTiles);
	for (size_t i = 0; i < n; ++i) {
		const float *s = samples + i * 6;
		const int px = (int) floorf(s[0]) - film.cropX, py = (int) floorf(s[1]) - film.cropY;
		REQUIRE(ctx, px >= 0 && py >= 0 && px < film.cropW && py < film.cropH, MTSGPU_ERR_ARG,
			"film_splat: sample %zu at (%g, %g) lies outside the crop window", i, s[0], s[1]);
		buckets[(py / TILE) * tilesX + px / TILE].push_back((uint32_t) i);
	}
	size_t maxCount = 1;
	for (auto &b : buckets) maxCount = std::max(maxCount, b.size());
	const uint32_t passes = (uint32_t) ((maxCount + TILE_PIX - 1) / TILE_PIX);
	const size_t perTile = (size_t) passes * TILE_PIX, total = perTile * nTiles;
	std::vector<float4> li(total, make_float4(0, 0, 0, 0));
	std::vector<float2> pos(total);
	float nanv; { uint32_t b = 0x7FC00000u; memcpy(&nanv, &b, 4); }
	for (auto &p : pos) p = make_float2(nanv, 0);
	for (uint32_t t = 0; t < nTiles; ++t)
		for (size_t k = 0; k < buckets[t].size(); ++k) {
			const float *s = samples + (size_t) buckets[t][k] * 6;
			li[t * perTile + k] = make_float4(s[2], s[3], s[4], s[5]);
			pos[t * perTile + k] = make_float2(s[0], s[1]);
		}
	const size_t filmFloats = (size_t) film.cropW * film.cropH * 5;
	DevBuf dTable;
	int rc;
	if ((rc = uploadFilter(ctx, dTable))) return rc;
	if ((rc = upload(ctx, ctx->dStage, li.data(), total * sizeof(float4)))) return rc;
	if ((rc = upload(ctx, ctx->dStageOut, pos.data(), total * sizeof(float2)))) return rc;
	CUDA_TRY(ctx, ctx->dFilm.alloc(filmFloats * 4));
	k_zero<<<blocksFor(filmFloats, 256), 256, 0, ctx->stream>>>(ctx->dFilm.as<float>(), filmFloats);
	BatchDesc b; memset(&b, 0, sizeof(b));
	b.tileBegin = 0; b.tileCount = nTiles; b.tilesX = tilesX; b.sampleCount = passes; b.slots = (uint32_t) total;
	const int smem = (TILE + 2 * film.border) * (TILE + 2 * film.border) * 5 * (int) sizeof(float);
	k_splat<<<nTiles, TILE_PIX, smem, ctx->stream>>>(film, b, ctx->dStage.as<float4>(), ctx->dStageOut.as<float2>(), dTable.as<float>(), ctx->dFilm.as<float>());
	return download(ctx, outFilm, ctx->dFilm, filmFloats * 4);
}
